// Row-streaming helpers and the single-setting linear / affine score kernels, shared by vx_align.cu (queries of up to 32
// tokens, and every other score kernel) and vx_align_long.cu (33 .. 64 tokens: a translation unit of its own, because the
// long instances take minutes to compile).
#pragma once
#include "vx_internal.cuh"

#include <math_constants.h>

template<int TP>
__device__ __forceinline__ void load_row(const float *__restrict__ s, int32_t id, float (&row)[TP]) {
	if constexpr (TP % 8 == 0) {
		// rows of 32 / 64 bytes: 256-bit loads (LDG.E.256, new on sm_100) halve the requests of this gather,
		// which is what bounds the linear / affine kernels (32 different rows per warp and token)
		const float *p = s + (size_t)id * TP;
		#pragma unroll
		for (int i = 0; i < TP / 8; i++) {
			asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
				: "=f"(row[8 * i + 0]), "=f"(row[8 * i + 1]), "=f"(row[8 * i + 2]), "=f"(row[8 * i + 3]),
				  "=f"(row[8 * i + 4]), "=f"(row[8 * i + 5]), "=f"(row[8 * i + 6]), "=f"(row[8 * i + 7])
				: "l"(p + 8 * i));
		}
	} else {
		const float4 *p = reinterpret_cast<const float4*>(s + (size_t)id * TP);
		#pragma unroll
		for (int i = 0; i < TP / 4; i++) {
			const float4 x = __ldg(p + i);
			row[4 * i + 0] = x.x;
			row[4 * i + 1] = x.y;
			row[4 * i + 2] = x.z;
			row[4 * i + 3] = x.w;
		}
	}
}

// Row of lane g of G lanes that share a token: the G lanes' rows are interleaved in chunks of one load instruction
// (CW floats), s[id][NF / CW][G][CW], so that every load instruction of the G lanes covers one contiguous G * CW * 4-byte
// piece -- with plain [id][G][NF] rows each instruction would touch every 128-byte line of the token's G * NF floats and
// the gather would cost NF / CW times the L1 wavefronts.  G = 1 is the plain row-major S of a single query.
template<int NF>
__device__ __forceinline__ void load_row_g(const float *__restrict__ s, int32_t id, int g_n, int g, float (&row)[NF]) {
	constexpr int CW = NF % 8 == 0 ? 8 : 4;
	constexpr int NI = NF / CW;
	const float *p = s + ((size_t)id * NI * g_n + g) * CW;
	#pragma unroll
	for (int i = 0; i < NI; i++) {
		const float *pi = p + (size_t)i * g_n * CW;
		if constexpr (CW == 8) {
			asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
				: "=f"(row[8 * i + 0]), "=f"(row[8 * i + 1]), "=f"(row[8 * i + 2]), "=f"(row[8 * i + 3]),
				  "=f"(row[8 * i + 4]), "=f"(row[8 * i + 5]), "=f"(row[8 * i + 6]), "=f"(row[8 * i + 7])
				: "l"(pi));
		} else {
			const float4 x = __ldg(reinterpret_cast<const float4*>(pi));
			row[4 * i + 0] = x.x;
			row[4 * i + 1] = x.y;
			row[4 * i + 2] = x.z;
			row[4 * i + 3] = x.w;
		}
	}
}

// The S rows of one window's tokens, one row ahead of their use: fn(row, u) is called for u = 0 .. len - 1 in order with
// the row of lane g of g_n (load_row_g).  Four token positions (one 16-byte quad of ids) per trip over two row buffers
// used in turn: no dynamic select on u & 3 and no register copies between a prefetch buffer and the row in use -- in the
// one-row-per-trip form those were more than half of the instructions of a DP row (scripts/sass_loop.py).
// Deeper prefetching was tried for the multi-query kernels (latency-bound on L2) and dropped: with four buffers ptxas
// moves the loads back towards their uses and shares scoreboards between them (scripts/sass_sb.py), and L1 prefetch
// requests (CCTL.E.PF1) cost more than they hide (profiles/r2_mq).
// The id stream may run up to two quads past the window (the next tile, or VX_TOK_SLACK); S rows are only loaded for
// u < len.
template<int NF, class Fn>
__device__ __forceinline__ void for_each_row(const float *__restrict__ s, const int32_t *tok_lane, int len, int g_n, int g, Fn &&fn) {
	const int4 *q = reinterpret_cast<const int4*>(tok_lane);
	int4 idq = __ldg(q);
	int4 idn = __ldg(q + 32);
	q += 64;
	int u = 0;
	float ra[NF], rb[NF];
	load_row_g<NF>(s, idq.x, g_n, g, ra);
	#pragma unroll 1
	for (; u + 4 <= len; u += 4) {
		load_row_g<NF>(s, idq.y, g_n, g, rb);
		fn(ra, u);
		load_row_g<NF>(s, idq.z, g_n, g, ra);
		fn(rb, u + 1);
		load_row_g<NF>(s, idq.w, g_n, g, rb);
		fn(ra, u + 2);
		idq = idn;
		idn = __ldg(q);
		q += 32;
		if (u + 4 < len) {
			load_row_g<NF>(s, idq.x, g_n, g, ra);
		}
		fn(rb, u + 3);
	}
	const int rem = len - u;
	if (rem > 0) {
		if (rem > 1) {
			load_row_g<NF>(s, idq.y, g_n, g, rb);
		}
		fn(ra, u);
		if (rem > 1) {
			if (rem > 2) {
				load_row_g<NF>(s, idq.z, g_n, g, ra);
			}
			fn(rb, u + 1);
			if (rem > 2) {
				fn(ra, u + 2);
			}
		}
	}
}

__device__ __forceinline__ float finish_score(float raw, float total, float boost) {
	// Score::Score (cpp/match/match.h:302-307) with reference_score == total for submatch_weight == 0.
	return (raw / total) * boost;
}

struct ScoreArgs {
	const VxTile *tiles;
	int64_t n_tiles;
	const int32_t *tok;
	const float *s;        // [V x TP]
	const float *boost;    // by slot or nullptr
	float *score;          // by slot
	int32_t t;
	float total;
	const int8_t *pos;     // tiled POS ids (tag-weighted similarity only)
	VxTagArgs tag;
};

// ---------------------------------------------------------------------------
// linear gap cost: Needleman-Wunsch / Smith-Waterman / semiglobal
// ---------------------------------------------------------------------------

template<int TP, int LOC>
__global__ void __launch_bounds__(128) score_linear_kernel(ScoreArgs a, float gs, float gt) {
	const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (tile >= a.n_tiles) {
		return;
	}
	const int lane = threadIdx.x & 31;
	const VxTile tl = a.tiles[tile];
	const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + lane) * 4;
	const int len = tl.len;
	const int t = a.t;

	float h[TP];   // h[v] = V[u-1, v]
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		h[v] = LOC == VX_GLOBAL ? -((float)(v + 1) * gt) : 0.0f;   // solver.h:2804-2806
	}
	float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
	float last_col = 0.0f;

	for_each_row<TP>(a.s, tok, len, 1, 0, [&](const float (&row)[TP], int u) {
		float cur[TP];
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			cur[v] = row[v];
		}
		if constexpr (TP <= VX_MAX_T_FAST) {   // tag weighting stops at 16 query tokens (make_plan)
			if (a.tag.on) {
				vx_apply_tag<TP>(a.tag, a.pos[vx_elem(tl.base128, u, lane)], cur);
			}
		}

		float diag = LOC == VX_GLOBAL ? -((float)u * gs) : 0.0f;        // V[u-1, -1]
		float left = LOC == VX_GLOBAL ? -((float)(u + 1) * gs) : 0.0f;  // V[u, -1]
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			const float up = h[v];
			float x = fmaxf(fmaxf(diag + cur[v], up - gs), left - gt);
			if (LOC == VX_LOCAL) {
				x = fmaxf(x, 0.0f);
				if (v < t) {
					best = fmaxf(best, x);
				}
			}
			diag = up;
			h[v] = x;
			left = x;
			if (v == t - 1) {
				last_col = x;
			}
		}
		if (LOC == VX_SEMIGLOBAL) {
			best = fmaxf(best, last_col);
		}
	});

	float raw;
	if (LOC == VX_LOCAL) {
		raw = best;
	} else if (LOC == VX_GLOBAL) {
		raw = last_col;
	} else {
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			if (v < t) {
				best = fmaxf(best, h[v]);
			}
		}
		raw = best;
	}

	const int64_t slot = tile * 32 + lane;
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	a.score[slot] = lane < tl.n_valid ? finish_score(raw, a.total, boost) : -CUDART_INF_F;
}

// ---------------------------------------------------------------------------
// affine gap cost: Gotoh
// ---------------------------------------------------------------------------

struct AffineArgs {
	float open_s, ext_s, open_t, ext_t;
};

__device__ __forceinline__ float affine_w(float ext, float open0, int k) {
	return k == 0 ? 0.0f : ext * (float)k + open0;   // AffineCost, solver.h:2850-2873
}

template<int TP, int LOC>
__global__ void __launch_bounds__(128) score_affine_kernel(ScoreArgs a, AffineArgs g) {
	const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (tile >= a.n_tiles) {
		return;
	}
	const int lane = threadIdx.x & 31;
	const VxTile tl = a.tiles[tile];
	const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + lane) * 4;
	const int len = tl.len;
	const int t = a.t;

	const float w1s = g.ext_s + g.open_s;
	const float w1t = g.ext_t + g.open_t;

	float d[TP], p[TP];   // D[u-1, v], P[u-1, v]
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		d[v] = LOC == VX_GLOBAL ? -affine_w(g.ext_t, g.open_t, v + 1) : 0.0f;
		p[v] = -CUDART_INF_F;
	}
	float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
	float last_col = 0.0f;

	for_each_row<TP>(a.s, tok, len, 1, 0, [&](const float (&row)[TP], int u) {
		float cur[TP];
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			cur[v] = row[v];
		}
		if constexpr (TP <= VX_MAX_T_FAST) {   // tag weighting stops at 16 query tokens (make_plan)
			if (a.tag.on) {
				vx_apply_tag<TP>(a.tag, a.pos[vx_elem(tl.base128, u, lane)], cur);
			}
		}

		float diag = LOC == VX_GLOBAL ? -affine_w(g.ext_s, g.open_s, u) : 0.0f;       // D[u-1, -1]
		float left = LOC == VX_GLOBAL ? -affine_w(g.ext_s, g.open_s, u + 1) : 0.0f;   // D[u, -1]
		float q = -CUDART_INF_F;                                                       // Q[u, -1]
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			const float up = d[v];
			const float pv = fmaxf(up - w1s, p[v] - g.ext_s);      // Gotoh (4)
			q = fmaxf(left - w1t, q - g.ext_t);                    // Gotoh (5)
			float x = fmaxf(fmaxf(diag + cur[v], pv), q);          // Gotoh (1)
			if (LOC == VX_LOCAL) {
				x = fmaxf(x, 0.0f);
				if (v < t) {
					best = fmaxf(best, x);
				}
			}
			p[v] = pv;
			diag = up;
			d[v] = x;
			left = x;
			if (v == t - 1) {
				last_col = x;
			}
		}
		if (LOC == VX_SEMIGLOBAL) {
			best = fmaxf(best, last_col);
		}
	});

	float raw;
	if (LOC == VX_LOCAL) {
		raw = best;
	} else if (LOC == VX_GLOBAL) {
		raw = last_col;
	} else {
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			if (v < t) {
				best = fmaxf(best, d[v]);
			}
		}
		raw = best;
	}

	const int64_t slot = tile * 32 + lane;
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	a.score[slot] = lane < tl.n_valid ? finish_score(raw, a.total, boost) : -CUDART_INF_F;
}

// queries of 33 .. 64 tokens on the linear / affine kernels (vx_align_long.cu); returns launches or < 0
int vx_launch_score_long(vx_corpus *c, const vx_sim *s, const ScoreArgs &a, const VxAlignDev &ap);
