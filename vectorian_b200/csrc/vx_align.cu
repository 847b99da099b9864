// Score-only alignment kernels (the hot loop).
//
// Replaces MatcherImpl::run_matches -> InjectiveAlignment::make_match -> {Linear,Affine,General}
// GapCostSolver::solve for every window of the corpus
// (vectorian/core/cpp/match/matcher_impl.h:71-109, cpp/metric/alignment.h:247-294,
//  pyalign/algorithm/solver.h:2817-2847, 2962-3028, 3130-3175).
//
// The reference fills a value + traceback matrix per window.  The maximum of a set of fp32
// candidates does not depend on the order they are compared in, so the SCORE needs no traceback and
// no tie-breaking: these kernels keep one DP row in registers and emit only the normalised score.
// Tracebacks are recomputed for the k winners by vx_trace.cu, which does reproduce the reference's
// candidate order.  fp32 add / compare only, compiled with -fmad=false: bit-identical to the oracle.
//
// Mapping: one warp per tile of 32 equal-length windows, one window per lane; the warp walks the
// token positions in lock step (no divergence) and every token-id load is one coalesced 128 B line.
#include "vx_align_rows.cuh"

// ---------------------------------------------------------------------------
// linear gap cost, several settings of one query at once (vx_search_batch, BASELINE config 5):
// NL Local settings followed by NG Global settings, each with its own gap costs.  The single-setting kernel is
// bound by the gather of the S rows (32 different rows per warp and token); here one gathered row feeds NL + NG
// dynamic-programming columns, so the work per gathered byte is NL + NG times higher and the kernel moves to
// the FP32 issue limit.  Cell values are the same floats as in score_linear_kernel (max of the same candidates).
// ---------------------------------------------------------------------------

constexpr int VX_MULTI_MAX = 8;

struct MultiArgs {
	float gs[VX_MULTI_MAX], gt[VX_MULTI_MAX];
	float *score[VX_MULTI_MAX];
};

__device__ __forceinline__ float fmax3(float a, float b, float c) {
	float r;
	asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
	return r;
}

// packed fp32 add (add.rn.f32x2 -> FADD2): two independent round-to-nearest adds in one issue slot
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
	float2 r;
	asm("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0, %1}, rc; }"
		: "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
	return r;
}

// Settings are processed in PAIRS (i, i + 1) of the same locality (NL and NG are even): the three adds of a cell --
// diag + S, up - g_s, left - g_t -- are three FADD2 over the pair, with {-g_s[i], -g_s[i+1]} / {-g_t[i], -g_t[i+1]} from the
// parameter bank (x - g == x + (-g) bit for bit); the maximum is one FMNMX3 per setting.  5 issue slots per cell and
// setting pair instead of 8.
template<int TP, int NL, int NG, bool ZC>
__global__ void __launch_bounds__(128) score_linear_multi_kernel(ScoreArgs a, const __grid_constant__ MultiArgs m) {
	constexpr int NS = NL + NG;
	constexpr int NP = NS / 2, NLP = NL / 2;
	static_assert(NL % 2 == 0 && NG % 2 == 0, "settings are paired");
	const int64_t tile = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (tile >= a.n_tiles) {
		return;
	}
	const int lane = threadIdx.x & 31;
	const VxTile tl = a.tiles[tile];
	const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + lane) * 4;
	const int len = tl.len;

	float2 h[NP][TP];                 // h[p][v] = {V_2p[u-1, v], V_2p+1[u-1, v]}
	float2 best[NLP > 0 ? NLP : 1];   // Local: running maxima (padded columns v >= t only repeat or lower values)
	#pragma unroll
	for (int p = 0; p < NP; p++) {
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			h[p][v] = p < NLP ? make_float2(0.0f, 0.0f)
				: make_float2(-((float)(v + 1) * m.gt[2 * p]), -((float)(v + 1) * m.gt[2 * p + 1]));   // solver.h:2804-2806
		}
		if (p < NLP) {
			best[p] = make_float2(0.0f, 0.0f);
		}
	}

	for_each_row<TP>(a.s, tok, len, 1, 0, [&](const float (&row)[TP], int u) {
		#pragma unroll
		for (int p = 0; p < NP; p++) {
			const float2 ngs = make_float2(-m.gs[2 * p], -m.gs[2 * p + 1]);
			const float2 ngt = make_float2(-m.gt[2 * p], -m.gt[2 * p + 1]);
			// V[u-1, -1] and V[u, -1]
			float2 diag = p < NLP ? make_float2(0.0f, 0.0f) : make_float2(-((float)u * m.gs[2 * p]), -((float)u * m.gs[2 * p + 1]));
			float2 left = p < NLP ? make_float2(0.0f, 0.0f) : make_float2(-((float)(u + 1) * m.gs[2 * p]), -((float)(u + 1) * m.gs[2 * p + 1]));
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				const float2 up = h[p][v];
				const float2 c0 = add2(diag, make_float2(row[v], row[v]));
				const float2 c1 = add2(up, ngs);
				const float2 c2 = add2(left, ngt);
				float2 x = make_float2(fmax3(c0.x, c1.x, c2.x), fmax3(c0.y, c1.y, c2.y));
				if (p < NLP && ZC) {
					x.x = fmaxf(x.x, 0.0f);   // Local zero; with S >= 0 it never wins: diag + S >= 0 already (ZC = false)
					x.y = fmaxf(x.y, 0.0f);
				}
				diag = up;
				h[p][v] = x;
				left = x;
			}
			if (p < NLP) {
				#pragma unroll
				for (int v = 0; v + 1 < TP; v += 2) {
					best[p].x = fmax3(best[p].x, h[p][v].x, h[p][v + 1].x);
					best[p].y = fmax3(best[p].y, h[p][v].y, h[p][v + 1].y);
				}
			}
		}
	});

	const int64_t slot = tile * 32 + lane;
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	#pragma unroll
	for (int i = 0; i < NS; i++) {
		const int p = i / 2;
		float raw;
		if (i < NL) {
			raw = (i & 1) ? best[p].y : best[p].x;
		} else {
			raw = (i & 1) ? h[p][0].y : h[p][0].x;   // V[len-1, t-1]; the selects are opaque to the compiler, which would
			#pragma unroll                           // otherwise turn the chain into a dynamically indexed local-memory array
			for (int v = 1; v < TP; v++) {
				asm("{ .reg .pred p; setp.eq.s32 p, %2, %3; selp.f32 %0, %1, %0, p; }"
					: "+f"(raw) : "f"((i & 1) ? h[p][v].y : h[p][v].x), "r"(v), "r"(a.t - 1));
			}
		}
		m.score[i][slot] = lane < tl.n_valid ? finish_score(raw, a.total, boost) : -CUDART_INF_F;
	}
}

// ---------------------------------------------------------------------------
// linear gap cost, several QUERIES at once (vx_search_batch with more than one query and one setting):
// the reference runs index.find once per query (vectorian/index.py:481-501).  The single-query kernel sits at the
// L1TEX wavefront limit: its 32 lanes gather 32 different S rows, so every load instruction touches 32 lines.
// Here the similarity matrices of K queries are interleaved into S_multi (K rows per token, chunked by load instruction:
// interleave_sim_kernel) and a lane is a (window, query) pair: lane = w * K + k.  Each load instruction of the K lanes of
// a window reads ONE contiguous K * 16- or 32-byte piece, so it touches 32 / K windows' pieces (1-2 lines each) instead of 32
// rows, and the per-lane DP state is as small as in the single-query kernel.  A warp takes 32 / K windows of a tile; the cells are the same fp32 maxima of the same candidates.
// ---------------------------------------------------------------------------

struct MqArgs {
	const float *s_multi;            // [V][chunk][lane][CW], see interleave_sim_kernel
	float gs, gt;
	float *score[VX_MULTI_MAX];      // per query, by slot
	int32_t t[VX_MULTI_MAX];
	float total[VX_MULTI_MAX];
};

template<int TP, int K, int LOC>
__global__ void __launch_bounds__(128) score_linear_mq_kernel(ScoreArgs a, const __grid_constant__ MqArgs m) {
	constexpr int W = 32 / K;        // windows per warp
	const int64_t unit = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int64_t tile = unit / K;
	const int sub = (int)(unit - tile * K);
	if (tile >= a.n_tiles) {
		return;
	}
	const int lane = threadIdx.x & 31;
	const int k = lane % K;
	const int wl = sub * W + lane / K;          // window (lane of the tile) this thread works on
	const VxTile tl = a.tiles[tile];
	const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + wl) * 4;
	const int len = tl.len;
	const int t = m.t[k];
	const float gs = m.gs, gt = m.gt;

	float h[TP];   // h[v] = V[u-1, v]
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		h[v] = LOC == VX_GLOBAL ? -((float)(v + 1) * gt) : 0.0f;
	}
	float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
	float last_col = 0.0f;

	for_each_row<TP>(m.s_multi, tok, len, K, k, [&](const float (&row)[TP], int u) {
		float diag = LOC == VX_GLOBAL ? -((float)u * gs) : 0.0f;
		float left = LOC == VX_GLOBAL ? -((float)(u + 1) * gs) : 0.0f;
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			const float up = h[v];
			float x = fmaxf(fmaxf(diag + row[v], up - gs), left - gt);
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
	const int64_t slot = tile * 32 + wl;
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	m.score[k][slot] = wl < tl.n_valid ? finish_score(raw, m.total[k], boost) : -CUDART_INF_F;
}

// Packed variant (Local / Global, non-negative gap costs): a lane is a (window, PAIR of queries).  S is interleaved as
// {S_a[v], S_b[v]} pairs (chunked like the rows of the kernel above), so a row load lands in register pairs and the three
// adds of a cell are three FADD2 over the pair -- {diag + S}, {up - g_s}, {left - g_t} -- followed by one FMNMX3 per query: 5 issue slots per
// cell PAIR (+ the running best) instead of 12.  The multi-query kernel above is issue-bound (84 % issue-active,
// profiles/r2_mq), so the instruction count is what matters.  Same fp32 operations on the same operands.
template<int TP, int KP, int LOC, bool ZC>
__global__ void __launch_bounds__(128, TP <= 8 ? 6 : TP <= 12 ? 4 : 3) score_linear_mq2_kernel(ScoreArgs a, const __grid_constant__ MqArgs m) {
	constexpr int W = 32 / KP;       // windows per warp; KP = pairs of queries per window
	const int64_t unit = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int64_t tile = unit / KP;
	const int sub = (int)(unit - tile * KP);
	if (tile >= a.n_tiles) {
		return;
	}
	const int lane = threadIdx.x & 31;
	const int p = lane % KP;
	const int wl = sub * W + lane / KP;
	const VxTile tl = a.tiles[tile];
	const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + wl) * 4;
	const int len = tl.len;
	const int ta = m.t[2 * p], tb = m.t[2 * p + 1];
	const float gs = m.gs, gt = m.gt;
	const float2 ngs = make_float2(-gs, -gs), ngt = make_float2(-gt, -gt);

	float2 h[TP];   // {V_a[u-1, v], V_b[u-1, v]}
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		const float b = LOC == VX_GLOBAL ? -((float)(v + 1) * gt) : 0.0f;
		h[v] = make_float2(b, b);
	}
	float2 best = make_float2(0.0f, 0.0f);

	for_each_row<2 * TP>(m.s_multi, tok, len, KP, p, [&](const float (&row)[2 * TP], int u) {
		const float d0 = LOC == VX_GLOBAL ? -((float)u * gs) : 0.0f;        // V[u-1, -1]
		const float l0 = LOC == VX_GLOBAL ? -((float)(u + 1) * gs) : 0.0f;  // V[u, -1]
		float2 diag = make_float2(d0, d0), left = make_float2(l0, l0);
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			const float2 up = h[v];
			const float2 c0 = add2(diag, make_float2(row[2 * v], row[2 * v + 1]));
			const float2 c1 = add2(up, ngs);
			const float2 c2 = add2(left, ngt);
			float2 x = make_float2(fmax3(c0.x, c1.x, c2.x), fmax3(c0.y, c1.y, c2.y));
			if (LOC == VX_LOCAL && ZC) {
				x.x = fmaxf(x.x, 0.0f);   // Local zero; with S >= 0 it never wins: diag + S >= 0 already (ZC = false)
				x.y = fmaxf(x.y, 0.0f);
			}
			diag = up;
			h[v] = x;
			left = x;
		}
		if (LOC == VX_LOCAL) {
			// padded columns (v >= t, S = 0) only repeat or lower earlier values (gap costs >= 0): the maximum over all
			// TP columns is the maximum over the query's own
			#pragma unroll
			for (int v = 0; v + 1 < TP; v += 2) {
				best.x = fmax3(best.x, h[v].x, h[v + 1].x);
				best.y = fmax3(best.y, h[v].y, h[v + 1].y);
			}
		}
	});

	float ra, rb;
	if (LOC == VX_LOCAL) {
		ra = best.x;
		rb = best.y;
	} else {
		ra = h[0].x;   // V[len-1, t-1] of either query; opaque selects keep the row in registers
		rb = h[0].y;
		#pragma unroll
		for (int v = 1; v < TP; v++) {
			asm("{ .reg .pred q; setp.eq.s32 q, %2, %3; selp.f32 %0, %1, %0, q; }" : "+f"(ra) : "f"(h[v].x), "r"(v), "r"(ta - 1));
			asm("{ .reg .pred q; setp.eq.s32 q, %2, %3; selp.f32 %0, %1, %0, q; }" : "+f"(rb) : "f"(h[v].y), "r"(v), "r"(tb - 1));
		}
	}
	const int64_t slot = tile * 32 + wl;
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	const bool valid = wl < tl.n_valid;
	m.score[2 * p][slot] = valid ? finish_score(ra, m.total[2 * p], boost) : -CUDART_INF_F;
	m.score[2 * p + 1][slot] = valid ? finish_score(rb, m.total[2 * p + 1], boost) : -CUDART_INF_F;
}

// S_multi: the K queries' similarity rows of one token, interleaved (below); zero beyond a query's own columns
struct InterleaveArgs {
	const float *s[VX_MULTI_MAX];
	int32_t tp[VX_MULTI_MAX];
};

// The layout load_row_g reads: out[row][chunk][lane g][CW].  Unpaired: lane g = query k, its row the tp columns of S_k
// (CW = 8 when tp % 8 == 0, else 4).  Paired (score_linear_mq2_kernel): lane g = queries 2g, 2g + 1, its row the 2 * tp floats
// {S_2g[j], S_2g+1[j]} for j = 0 .. tp - 1 (CW = 8).
__global__ void __launch_bounds__(256) interleave_sim_kernel(const __grid_constant__ InterleaveArgs g, int64_t rows, int k_n, int tp,
	float *__restrict__ out, bool paired) {

	const int64_t n = rows * k_n * tp;
	const int nf = paired ? 2 * tp : tp;
	const int cw = nf % 8 == 0 ? 8 : 4;
	const int ni = nf / cw;
	const int g_n = paired ? k_n / 2 : k_n;
	for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
		const int c = (int)(i % cw);
		int64_t q = i / cw;
		const int lane = (int)(q % g_n);
		q /= g_n;
		const int chunk = (int)(q % ni);
		const int64_t r = q / ni;
		const int e = chunk * cw + c;   // element of the lane's row
		const int j = paired ? e >> 1 : e;
		const int k = paired ? 2 * lane + (e & 1) : lane;
		out[i] = j < g.tp[k] ? g.s[k][r * g.tp[k] + j] : 0.0f;
	}
}

// ---------------------------------------------------------------------------
// general gap cost: Waterman-Smith-Beyer, any window length (column history in global scratch)
// ---------------------------------------------------------------------------
//
// V[u,v] = max( [0], V[u-1,v-1] + sim, max_{k<u} V[k,v] - w_s[u-k], max_{k<v} V[u,k] - w_t[v-k] )
// (solver.h:3145-3173).  Persistent warps stride over the tiles; each warp owns a scratch region
// hist[u][v][lane] holding the finished rows of its current tile.

template<int TP, int LOC>
__global__ void __launch_bounds__(128) score_general_scratch_kernel(
	ScoreArgs a, const float *__restrict__ w_s, const float *__restrict__ w_t,
	float *__restrict__ scratch, int32_t max_len) {

	const int lane = threadIdx.x & 31;
	const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const int64_t n_warps = (int64_t)gridDim.x * (blockDim.x >> 5);
	float *hist = scratch + (size_t)warp * max_len * TP * 32 + lane;
	const int t = a.t;

	float wt[TP + 1];
	#pragma unroll
	for (int k = 0; k <= TP; k++) {
		wt[k] = k <= t ? __ldg(w_t + k) : 0.0f;
	}

	for (int64_t tile = warp; tile < a.n_tiles; tile += n_warps) {
		const VxTile tl = a.tiles[tile];
		const int32_t *tok = a.tok + ((size_t)tl.base128 * 32 + lane) * 4;
		const int len = tl.len;

		float prev[TP];  // V[u-1, v]; V[-1, v] = the top border
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			prev[v] = LOC == VX_GLOBAL ? -wt[v + 1] : 0.0f;
		}
		float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
		float last_col = 0.0f;

		// row u from its vertical candidates (k = -1 .. u-1 folded into vert), the diagonal and the horizontal scan;
		// leaves the row in prev and in the history
		auto finish_row = [&](const float (&vert)[TP], const float (&sim)[TP], int u) {
			const float left_border = LOC == VX_GLOBAL ? -__ldg(w_s + u + 1) : 0.0f;   // V[u, -1]
			float diag = LOC == VX_GLOBAL ? (u == 0 ? -wt[0] : -__ldg(w_s + u)) : 0.0f;  // V[u-1, -1]
			float cur[TP];
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				float x = fmaxf(diag + sim[v], vert[v]);
				x = fmaxf(x, left_border - wt[v + 1]);
				#pragma unroll
				for (int k = 0; k < v; k++) {
					x = fmaxf(x, cur[k] - wt[v - k]);
				}
				if (LOC == VX_LOCAL) {
					x = fmaxf(x, 0.0f);
					if (v < t) {
						best = fmaxf(best, x);
					}
				}
				diag = prev[v];
				cur[v] = x;
				if (v == t - 1) {
					last_col = x;
				}
			}
			float *hu = hist + (size_t)u * TP * 32;
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				prev[v] = cur[v];
				hu[v * 32] = cur[v];
			}
			if (LOC == VX_SEMIGLOBAL) {
				best = fmaxf(best, last_col);
			}
		};

		// U rows per pass over the history: every finished row k < u is read once for rows u .. u + U - 1 (the history lives
		// in L2 and its traffic is what bounds this kernel); a row of the block is folded into the later rows' vertical
		// candidates from registers as soon as it is finished.  U = 4 up to 24 query columns, 2 beyond (registers).
		constexpr int U = TP <= 24 ? 4 : 2;
		int u = 0;
		for (; u + U <= len; u += U) {
			float vert[U][TP];
			#pragma unroll
			for (int i = 0; i < U; i++) {
				const float w = __ldg(w_s + u + i + 1);
				#pragma unroll
				for (int v = 0; v < TP; v++) {
					vert[i][v] = (LOC == VX_GLOBAL ? -wt[v + 1] : 0.0f) - w;   // V[-1, v] - w_s[u + i + 1]
				}
			}
			for (int k = 0; k < u; k++) {
				float wk[U];
				#pragma unroll
				for (int i = 0; i < U; i++) {
					wk[i] = __ldg(w_s + (u + i - k));
				}
				const float *hk = hist + (size_t)k * TP * 32;
				#pragma unroll
				for (int v = 0; v < TP; v++) {
					const float h = hk[v * 32];
					#pragma unroll
					for (int i = 0; i < U; i++) {
						vert[i][v] = fmaxf(vert[i][v], h - wk[i]);
					}
				}
			}
			#pragma unroll
			for (int i = 0; i < U; i++) {
				float sim[TP];
				load_row<TP>(a.s, tok[vx_elem(0, u + i, 0)], sim);
				finish_row(vert[i], sim, u + i);
				#pragma unroll
				for (int j = i + 1; j < U; j++) {
					const float w = __ldg(w_s + (j - i));
					#pragma unroll
					for (int v = 0; v < TP; v++) {
						vert[j][v] = fmaxf(vert[j][v], prev[v] - w);   // k = u + i, from registers
					}
				}
			}
		}
		for (; u < len; u++) {
			float vert[TP];
			const float w_top = __ldg(w_s + u + 1);
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				vert[v] = (LOC == VX_GLOBAL ? -wt[v + 1] : 0.0f) - w_top;
			}
			for (int k = 0; k < u; k++) {
				const float wk = __ldg(w_s + (u - k));
				const float *hk = hist + (size_t)k * TP * 32;
				#pragma unroll
				for (int v = 0; v < TP; v++) {
					vert[v] = fmaxf(vert[v], hk[v * 32] - wk);
				}
			}
			float sim[TP];
			load_row<TP>(a.s, tok[vx_elem(0, u, 0)], sim);
			finish_row(vert, sim, u);
		}

		float raw;
		if (LOC == VX_LOCAL) {
			raw = best;
		} else if (LOC == VX_GLOBAL) {
			raw = last_col;
		} else {
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				if (v < t) {
					best = fmaxf(best, prev[v]);
				}
			}
			raw = best;
		}
		const int64_t slot = tile * 32 + lane;
		const float boost = a.boost ? a.boost[slot] : 1.0f;
		a.score[slot] = lane < tl.n_valid ? finish_score(raw, a.total, boost) : -CUDART_INF_F;
		__syncwarp();
	}
}

// ---------------------------------------------------------------------------
// dispatch
// ---------------------------------------------------------------------------

static int g_sm_count = 0;

static int sm_count() {
	if (g_sm_count == 0) {
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
	}
	return g_sm_count;
}

template<int TP, int LOC>
static int launch_score_tl(vx_corpus *c, const vx_sim *s, const ScoreArgs &a, const VxAlignDev &ap, const VxQueryDev &qp) {
	const int warps = 4;
	const unsigned grid = (unsigned)((a.n_tiles + warps - 1) / warps);
	switch (ap.gap_kind) {
		case VX_GAP_LINEAR:
			score_linear_kernel<TP, LOC><<<grid, warps * 32, 0, c->stream>>>(a, ap.lin_s, ap.lin_t);
			return 1;
		case VX_GAP_AFFINE: {
			AffineArgs g{ap.aff_open_s, ap.aff_ext_s, ap.aff_open_t, ap.aff_ext_t};
			score_affine_kernel<TP, LOC><<<grid, warps * 32, 0, c->stream>>>(a, g);
			return 1;
		}
		default: {
			// windows up to 64 tokens: shared-memory kernel (vx_wsb.cu); the rest: global scratch
			int64_t first = 0;
			int n_fast = 0;
			if (TP <= VX_MAX_T_FAST) {
				n_fast = vx_launch_wsb(c, s, ap, qp, &first, a.tok);   // (tag weighting: windows <= 64 tokens only, see make_plan)
			} else {
				first = c->n_tiles;    // queries of 17 .. 32 tokens: every tile goes to the scratch kernel below
			}
			if (n_fast < 0) {
				return -1;
			}
			if (first == 0) {
				return n_fast;
			}
			ScoreArgs a_long = a;
			a_long.n_tiles = first;
			// persistent warps: 16 per SM; 12 when every tile comes here (queries of 17 .. 32 tokens: 168 registers, and the
			// column histories in flight -- warps x window length x TP x 128 bytes -- should stay inside L2): 2 M sentences at
			// T = 24 take 5.65 ms with 12 warps per SM, 6.2 with 8, 6.4 with 16 (which do not all fit), 7.6 with 4
			const int blocks = sm_count() * (TP > VX_MAX_T_FAST ? 3 : 4);
			const size_t need = sizeof(float) * (size_t)blocks * warps * c->max_len * TP * 32;
			if (c->scratch_bytes < need) {
				if (c->d_scratch) {
					cudaFree(c->d_scratch);
					c->d_scratch = nullptr;
					c->scratch_bytes = 0;
				}
				if (cudaMalloc(&c->d_scratch, need) != cudaSuccess) {
					vx_set_error("cudaMalloc of the general-gap scratch failed");
					return -1;
				}
				c->scratch_bytes = need;
			}
			score_general_scratch_kernel<TP, LOC><<<blocks, warps * 32, 0, c->stream>>>(
				a_long, ap.w_s, ap.w_t, (float*)c->d_scratch, c->max_len);
			return n_fast + 1;
		}
	}
}

template<int TP>
static int launch_score_t(vx_corpus *c, const vx_sim *s, const ScoreArgs &a, const VxAlignDev &ap, const VxQueryDev &qp) {
	switch (ap.locality) {
		case VX_LOCAL:
			return launch_score_tl<TP, VX_LOCAL>(c, s, a, ap, qp);
		case VX_GLOBAL:
			return launch_score_tl<TP, VX_GLOBAL>(c, s, a, ap, qp);
		default:
			return launch_score_tl<TP, VX_SEMIGLOBAL>(c, s, a, ap, qp);
	}
}

template<int TP>
static int launch_multi_t(const ScoreArgs &a, const MultiArgs &m, int nl, int ng, bool zc, cudaStream_t st) {
	const int warps = 4;
	const unsigned grid = (unsigned)((a.n_tiles + warps - 1) / warps);
	#define VX_MULTI_CASE(NL, NG) if (nl == NL && ng == NG) { \
		if (zc && NL > 0) { \
			score_linear_multi_kernel<TP, NL, NG, true><<<grid, warps * 32, 0, st>>>(a, m); \
		} else { \
			score_linear_multi_kernel<TP, NL, NG, false><<<grid, warps * 32, 0, st>>>(a, m); \
		} \
		return 1; \
	}
	VX_MULTI_CASE(4, 4)
	VX_MULTI_CASE(4, 0)
	VX_MULTI_CASE(0, 4)
	VX_MULTI_CASE(2, 2)
	VX_MULTI_CASE(2, 0)
	VX_MULTI_CASE(0, 2)
	#undef VX_MULTI_CASE
	return 0;
}

// Several linear-gap settings of one query in one pass over the corpus: nl Local settings then ng Global ones
// (gs / gt / score in that order).  Returns 1 when launched, 0 when this (nl, ng, tp) shape has no kernel.
int vx_launch_score_multi(vx_corpus *c, const vx_sim *s, const VxQueryDev &qp, int nl, int ng,
	const float *gs, const float *gt, float *const *score) {

	if (c->n_tiles == 0 || nl + ng > VX_MULTI_MAX) {
		return 0;
	}
	ScoreArgs a;
	a.tiles = c->d_tiles;
	a.n_tiles = c->n_tiles;
	a.tok = c->d_tok;
	a.s = s->d_s;
	a.boost = c->d_boost;
	a.score = nullptr;
	a.t = qp.t;
	a.total = qp.total;
	a.pos = nullptr;
	memset(&a.tag, 0, sizeof(a.tag));   // the multi-setting kernel has no tag weighting: the caller keeps such settings out
	MultiArgs m;
	memset(&m, 0, sizeof(m));
	for (int i = 0; i < nl + ng; i++) {
		m.gs[i] = gs[i];
		m.gt[i] = gt[i];
		m.score[i] = score[i];
	}
	switch (s->tp) {
		case 4:
			return launch_multi_t<4>(a, m, nl, ng, !s->nonneg, c->stream);
		case 8:
			return launch_multi_t<8>(a, m, nl, ng, !s->nonneg, c->stream);
		case 12:
			return launch_multi_t<12>(a, m, nl, ng, !s->nonneg, c->stream);
		case 16:
			return launch_multi_t<16>(a, m, nl, ng, !s->nonneg, c->stream);
		default:
			return 0;
	}
}

// Fast path preconditions: static similarity, T <= 16, plain (isolated) scoring with
// submatch_weight == 0, so that reference_score == T exactly.  Everything else is scored by the
// exact kernel in vx_trace.cu (the caller decides).
int vx_launch_score(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, const int32_t *ids) {
	if (c->n_tiles == 0) {
		return 0;
	}
	ScoreArgs a;
	a.tiles = c->d_tiles;
	a.n_tiles = c->n_tiles;
	a.tok = ids ? ids : c->d_tok;
	a.s = s->d_s;
	a.boost = c->d_boost;
	a.score = c->d_score;
	a.t = qp.t;
	a.total = qp.total;
	a.pos = c->d_pos;
	a.tag = vx_make_tag_args(qp);
	int n;
	switch (s->tp) {
		case 4:
			n = launch_score_t<4>(c, s, a, ap, qp);
			break;
		case 8:
			n = launch_score_t<8>(c, s, a, ap, qp);
			break;
		case 12:
			n = launch_score_t<12>(c, s, a, ap, qp);
			break;
		case 16:
			n = launch_score_t<16>(c, s, a, ap, qp);
			break;
		case 20:
			n = launch_score_t<20>(c, s, a, ap, qp);
			break;
		case 24:
			n = launch_score_t<24>(c, s, a, ap, qp);
			break;
		case 28:
			n = launch_score_t<28>(c, s, a, ap, qp);
			break;
		case 32:
			n = launch_score_t<32>(c, s, a, ap, qp);
			break;
		case 40:
		case 48:
		case 56:
		case 64:
			n = vx_launch_score_long(c, s, a, ap);   // vx_align_long.cu
			break;
		default:
			vx_set_error("vx_launch_score: query longer than the fast path supports");
			return -1;
	}
	if (n >= 0 && cudaGetLastError() != cudaSuccess) {
		vx_set_error("score kernel launch failed");
		return -1;
	}
	return n;
}

template<int TP, int K>
static int launch_mq_tk(const ScoreArgs &a, const MqArgs &m, int locality, cudaStream_t st) {
	const int warps = 4;
	const unsigned grid = (unsigned)((a.n_tiles * K + warps - 1) / warps);
	switch (locality) {
		case VX_LOCAL:
			score_linear_mq_kernel<TP, K, VX_LOCAL><<<grid, warps * 32, 0, st>>>(a, m);
			return 1;
		case VX_GLOBAL:
			score_linear_mq_kernel<TP, K, VX_GLOBAL><<<grid, warps * 32, 0, st>>>(a, m);
			return 1;
		default:
			score_linear_mq_kernel<TP, K, VX_SEMIGLOBAL><<<grid, warps * 32, 0, st>>>(a, m);
			return 1;
	}
}

template<int TP, int KP>
static int launch_mq2_tk(const ScoreArgs &a, const MqArgs &m, int locality, bool zc, cudaStream_t st) {
	const int warps = 4;
	const unsigned grid = (unsigned)((a.n_tiles * KP + warps - 1) / warps);
	if (locality == VX_LOCAL) {
		if (zc) {
			score_linear_mq2_kernel<TP, KP, VX_LOCAL, true><<<grid, warps * 32, 0, st>>>(a, m);
		} else {
			score_linear_mq2_kernel<TP, KP, VX_LOCAL, false><<<grid, warps * 32, 0, st>>>(a, m);
		}
	} else {
		score_linear_mq2_kernel<TP, KP, VX_GLOBAL, false><<<grid, warps * 32, 0, st>>>(a, m);
	}
	return 1;
}

template<int TP>
static int launch_mq2_t(const ScoreArgs &a, const MqArgs &m, int k_n, int locality, bool zc, cudaStream_t st) {
	switch (k_n) {
		case 8:
			return launch_mq2_tk<TP, 4>(a, m, locality, zc, st);
		case 4:
			return launch_mq2_tk<TP, 2>(a, m, locality, zc, st);
		case 2:
			return launch_mq2_tk<TP, 1>(a, m, locality, zc, st);
		default:
			return 0;
	}
}

template<int TP>
static int launch_mq_t(const ScoreArgs &a, const MqArgs &m, int k_n, int locality, cudaStream_t st) {
	switch (k_n) {
		case 8:
			return launch_mq_tk<TP, 8>(a, m, locality, st);
		case 4:
			return launch_mq_tk<TP, 4>(a, m, locality, st);
		case 2:
			return launch_mq_tk<TP, 2>(a, m, locality, st);
		default:
			return 0;
	}
}

// K = 2, 4 or 8 queries (static similarity matrices over the same vocabulary) against ONE linear-gap setting in one
// pass over the corpus: interleaves their S into d_s_multi [rows x K x tp] and scores every (window, query) pair.
// Returns launches (2) or 0 when the shape has no kernel.
int vx_launch_score_mq(vx_corpus *c, const vx_sim *const *sims, int k_n, int locality, float gs, float gt,
	float *d_s_multi, float *const *score) {

	if (c->n_tiles == 0 || (k_n != 2 && k_n != 4 && k_n != 8)) {
		return 0;
	}
	int tp = 4;
	InterleaveArgs g;
	memset(&g, 0, sizeof(g));
	MqArgs m;
	memset(&m, 0, sizeof(m));
	for (int k = 0; k < k_n; k++) {
		tp = sims[k]->tp > tp ? sims[k]->tp : tp;
		g.s[k] = sims[k]->d_s;
		g.tp[k] = sims[k]->tp;
		m.score[k] = score[k];
		m.t[k] = sims[k]->t;
		m.total[k] = (float)sims[k]->t;
	}
	const int64_t rows = sims[0]->rows;
	if (tp > VX_MAX_T_FAST || rows * k_n >= ((int64_t)1 << 31)) {
		return 0;
	}
	m.s_multi = d_s_multi;
	m.gs = gs;
	m.gt = gt;
	const int64_t n = rows * k_n * tp;
	int64_t blocks = (n + 255) / 256;
	blocks = blocks > 148 * 16 ? 148 * 16 : blocks < 1 ? 1 : blocks;
	// packed pairs of queries (FADD2) for Local / Global with non-negative gap costs; VX_MQ_NO_PACK: A/B switch
	static const bool no_pack = getenv("VX_MQ_NO_PACK") != nullptr;
	const bool paired = !no_pack && (locality == VX_LOCAL || locality == VX_GLOBAL) && gs >= 0.0f && gt >= 0.0f;
	bool zc = false;
	for (int k = 0; k < k_n; k++) {
		zc = zc || !sims[k]->nonneg;
	}
	interleave_sim_kernel<<<(unsigned)blocks, 256, 0, c->stream>>>(g, rows, k_n, tp, d_s_multi, paired);
	ScoreArgs a;
	a.tiles = c->d_tiles;
	a.n_tiles = c->n_tiles;
	a.tok = c->d_tok;
	a.s = nullptr;
	a.boost = c->d_boost;
	a.score = nullptr;
	a.t = 0;
	a.total = 0.0f;
	a.pos = nullptr;
	memset(&a.tag, 0, sizeof(a.tag));
	int launched;
	switch (tp) {
		case 4:
			launched = paired ? launch_mq2_t<4>(a, m, k_n, locality, zc, c->stream) : launch_mq_t<4>(a, m, k_n, locality, c->stream);
			break;
		case 8:
			launched = paired ? launch_mq2_t<8>(a, m, k_n, locality, zc, c->stream) : launch_mq_t<8>(a, m, k_n, locality, c->stream);
			break;
		case 12:
			launched = paired ? launch_mq2_t<12>(a, m, k_n, locality, zc, c->stream) : launch_mq_t<12>(a, m, k_n, locality, c->stream);
			break;
		case 16:
			launched = paired ? launch_mq2_t<16>(a, m, k_n, locality, zc, c->stream) : launch_mq_t<16>(a, m, k_n, locality, c->stream);
			break;
		default:
			launched = 0;
	}
	if (launched && cudaGetLastError() != cudaSuccess) {
		vx_set_error("multi-query score kernel launch failed");
		return -1;
	}
	return launched ? launched + 1 : 0;
}
