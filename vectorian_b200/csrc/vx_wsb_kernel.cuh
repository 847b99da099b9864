// Waterman-Smith-Beyer (general gap cost) score kernel: the roofline-critical kernel of the path.
//
// Replaces GeneralGapCostSolver::solve (pyalign/algorithm/solver.h:3130-3175) + the per-sentence
// loop of MatcherImpl::run_matches (vectorian/core/cpp/match/matcher_impl.h:71-109) for every window
// of at most VX_WSB_MAX_LEN tokens; longer windows take score_general_scratch_kernel (vx_align.cu).
//
//   V[u,v] = max( [0], V[u-1,v-1] + sim(u,v), max_{k<u} V[k,v] - w_s[u-k], max_{k<v} V[u,k] - w_t[v-k] )
//
// Work per window of L tokens against T query tokens: L*T*(L+T)/2 candidates, each one fp32 add and
// one fp32 max -- the kernel is bound by the fma/alu issue pipes, not by HBM (100 B of token ids per
// window).  Design (measured on B200 with scripts/ubench/pipes.cu: FADD r,c = 1 cyc/warp/SMSP,
// FMNMX3 = 2 cyc, so 2 candidates cost 3 issue slots when both adds read one register + one
// constant/uniform operand and the two maxima fold into one 3-input FMNMX3):
//   * one window per lane, 32 equal-length windows per warp (a tile): no divergence, token ids are
//     read as one 128-byte line per token position;
//   * rows are processed in blocks of U: the U x T accumulators of the block live in registers; a
//     finished row is read back from shared memory once per BLOCK (not once per row), so every
//     LDS.128 feeds 4*U candidates;
//   * finished rows (the column history the vertical gap scan needs) live in TENSOR MEMORY first:
//     the kernel allocates the SM's 512 TMEM columns and uses them as a per-thread scratchpad
//     (tcgen05.st / tcgen05.ld .32x32b: thread = TMEM lane, row u = T consecutive columns), 20 rows
//     per window at 8 warps per SM; rows beyond that go to shared memory as [u][v/4][lane] float4
//     (conflict-free LDS.128/STS.128).  TMEM + smem together hold 8 warps x 32 windows x 36 rows, which
//     is what lets two warps share each scheduler (one warp alone issues every other cycle at best);
//   * gap tables come from the kernel parameter bank (constant operands / uniform loads), the
//     similarity rows S[id, 0..T) from L2/L1 with float4 loads issued one block ahead of their use;
//   * persistent warps pull tiles from an atomic counter; one launch per length class so that the
//     shared-memory slice of a warp (and therefore warps per SM) matches the class.
// fp32 add / max only, compiled with -fmad=false: scores are bit-identical to the oracle.
#pragma once
#include "vx_internal.cuh"

#include <math_constants.h>
#include <cstring>

constexpr int VX_WSB_U = 4;

struct WsbArgs {
	const VxTile *tiles;
	const int32_t *tok;
	const float *s;          // [V x TP]
	const float *boost;      // by slot or nullptr
	float *score;            // by slot
	unsigned int *counter;   // chunk counter of this launch (zeroed by the host)
	int32_t tile_begin, tile_end;
	int32_t chunk;           // tiles per reservation for the first n_big reservations, then one tile each
	uint32_t n_big;
	int32_t t;
	int32_t cap;             // rows of history per warp in shared memory (rows rt .. rt+cap-1)
	// Long queue of the main launch: warps [0, n_long_warps) of every CTA own cap_long >= cap rows and first take the
	// tiles [long_begin, long_end) -- the next longer length classes -- one at a time, then join the others.
	int32_t n_long_warps, cap_long;
	int32_t long_begin;
	uint32_t n_long_tiles;
	unsigned int *counter_long;
	int32_t rt;              // rows [0, rt) of the history live in tensor memory (multiple of U)
	int32_t tmem_cols;       // TMEM columns owned by one warp (256: two warps per lane quarter, 512: one)
	float total;
	float ws[VX_WSB_MAX_LEN + 4];   // w_s[0..160], gap cost along the sentence
	float wt[VX_MAX_T_FAST + 4];    // w_t[0..16], gap cost along the query
	// the same tables negated and paired for the packed adds (add.rn.f32x2 -> FADD2, one issue slot for two
	// candidates; x - w == x + (-w) bit for bit):
	float2 nws2[VX_WSB_MAX_LEN + 4];   // {-w_s[d], -w_s[d]}: two neighbouring columns of one finished row
	float2 nwt2[VX_MAX_T_FAST + 4];    // {-w_t[d], -w_t[d-1]}: two neighbouring cells (k, k+1) left of column k + d
	const int8_t *pos;                 // tiled POS ids, same element order as tok (tag-weighted similarity only)
	VxTagArgs tag;
};

__device__ __forceinline__ float max3(float a, float b, float c) {
	float r;
	asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
	return r;
}

// Packed fp32 add (sm_100: add.rn.f32x2 -> SASS FADD2 Rd, Ra.F32x2, URb.F32x2): both halves are ordinary
// round-to-nearest adds.  The register pairs are formed by the allocator (the tcgen05.ld / LDS.128
// destinations and the accumulator rows are consecutive registers), the gap-cost pair comes from the
// parameter bank through a uniform register pair.
__device__ __forceinline__ float2 add2(float x0, float x1, float2 w) {
	float2 r;
	asm("{ .reg .b64 ra, rb, rc; mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; add.rn.f32x2 rc, ra, rb; mov.b64 {%0, %1}, rc; }"
		: "=f"(r.x), "=f"(r.y) : "f"(x0), "f"(x1), "f"(w.x), "f"(w.y));
	return r;
}

// ---- tensor memory as a per-thread scratchpad (SASS: LDTM / STTM) ----
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float *r) {
	uint32_t a, b, c, d;
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
		: "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr));
	r[0] = __uint_as_float(a);
	r[1] = __uint_as_float(b);
	r[2] = __uint_as_float(c);
	r[3] = __uint_as_float(d);
}

__device__ __forceinline__ void tmem_st4(uint32_t taddr, const float *r) {
	asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
		:: "r"(taddr), "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])),
		"r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])) : "memory");
}

__device__ __forceinline__ void tmem_wait_ld() {
	asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_wait_st() {
	asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// Running maximum that folds two candidates per FMNMX3.  After full unrolling `has` is a
// compile-time constant, so this costs ceil(n / 2) max instructions for n candidates.
struct Fold {
	float x, p;
	bool has;
	__device__ __forceinline__ explicit Fold(float x0) : x(x0), p(0.0f), has(false) {}
	__device__ __forceinline__ void push(float c) {
		if (has) {
			x = max3(x, p, c);
			has = false;
		} else {
			p = c;
			has = true;
		}
	}
	__device__ __forceinline__ void push2(float c0, float c1) {
		if (has) {
			x = max3(x, p, c0);
			p = c1;
		} else {
			x = max3(x, c0, c1);
		}
	}
	__device__ __forceinline__ float done() const {
		return has ? fmaxf(x, p) : x;
	}
};

template<int TP>
__device__ __forceinline__ void load_sim_row(const float4 *__restrict__ s4, int32_t id, float (&row)[TP]) {
	#pragma unroll
	for (int q = 0; q < TP / 4; q++) {
		const float4 x = __ldg(s4 + (size_t)id * (TP / 4) + q);
		row[4 * q + 0] = x.x;
		row[4 * q + 1] = x.y;
		row[4 * q + 2] = x.z;
		row[4 * q + 3] = x.w;
	}
}

__device__ __forceinline__ int32_t quad_get(const int4 &q, int i) {
	return i == 0 ? q.x : i == 1 ? q.y : i == 2 ? q.z : q.w;
}

// rows k, k+1 of the column history from tensor memory (asynchronous until tmem_wait_ld) / shared memory
template<int TP>
__device__ __forceinline__ void tmem_ld_pair(const uint32_t tmem, const int k, float (&h0)[TP], float (&h1)[TP]) {
	#pragma unroll
	for (int q = 0; q < TP / 4; q++) {
		tmem_ld4(tmem + (uint32_t)(k * TP + 4 * q), h0 + 4 * q);
		tmem_ld4(tmem + (uint32_t)((k + 1) * TP + 4 * q), h1 + 4 * q);
	}
}

template<int TP>
__device__ __forceinline__ void smem_ld_pair(const float4 *__restrict__ hist, const int k, float (&h0)[TP], float (&h1)[TP]) {
	constexpr int Q = TP / 4;
	#pragma unroll
	for (int q = 0; q < Q; q++) {
		const float4 x = hist[((size_t)k * Q + q) * 32];
		const float4 y = hist[((size_t)(k + 1) * Q + q) * 32];
		h0[4 * q + 0] = x.x; h0[4 * q + 1] = x.y; h0[4 * q + 2] = x.z; h0[4 * q + 3] = x.w;
		h1[4 * q + 0] = y.x; h1[4 * q + 1] = y.y; h1[4 * q + 2] = y.z; h1[4 * q + 3] = y.w;
	}
}

// Vertical candidates of the NR rows of a block from two finished rows k, k+1 (h0, h1); d = u0 - k.
//   acc[i][v] = max(acc[i][v], h0[v] - w_s[d + i], h1[v] - w_s[d + i - 1])
// F2: one FADD2 forms the candidates of two neighbouring columns, so 4 candidates cost 2 FADD2 + 2 FMNMX3
// (4 issue slots instead of 6).
template<int TP, int NR, bool F2>
__device__ __forceinline__ void wsb_history_pair(const WsbArgs &a, const int d, float (&acc)[NR][TP],
	const float (&h0)[TP], const float (&h1)[TP]) {

	#pragma unroll
	for (int i = 0; i < NR; i++) {
		if (F2) {
			const float2 w0 = a.nws2[d + i];
			const float2 w1 = a.nws2[d + i - 1];
			#pragma unroll
			for (int v = 0; v < TP; v += 2) {
				const float2 p = add2(h0[v], h0[v + 1], w0);
				const float2 q = add2(h1[v], h1[v + 1], w1);
				acc[i][v] = max3(acc[i][v], p.x, q.x);
				acc[i][v + 1] = max3(acc[i][v + 1], p.y, q.y);
			}
		} else {
			const float w0 = a.ws[d + i];
			const float w1 = a.ws[d + i - 1];
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				acc[i][v] = max3(acc[i][v], h0[v] - w0, h1[v] - w1);
			}
		}
	}
}

// One block of NR rows u0 .. u0+NR-1 of one tile.
//   prev[v]  in: V[u0-1, v]  out: V[u0+NR-1, v]
//   sim      in: S rows of this block; out: S rows of the NEXT block of the warp's token stream
//            (each row is re-loaded as soon as it has been consumed, ids from idq_next)
// TAG: tag-weighted similarity.  `posq` holds the POS bytes of this block's four tokens (applied to the S rows when
// they are consumed, so that the prefetch of the rows is not disturbed); the POS quad of the next block is requested
// through `pquads`, which runs one block ahead, and handed back in `posq` when the block is done.
template<int TP, int LOC, bool BORD, int NR, bool F2, bool PIPE, bool TAG>
__device__ __forceinline__ void wsb_block(
	const WsbArgs &a, const uint32_t tmem, float4 *__restrict__ hist, const int u0, const int t,
	float (&prev)[TP], float &best, float &last_col,
	float (&sim)[VX_WSB_U][TP], int4 &idq_next, const int4 *__restrict__ &quads,
	uint32_t &posq, const uint32_t *__restrict__ &pquads) {

	constexpr int Q = TP / 4;
	constexpr int U = VX_WSB_U;
	const float4 *__restrict__ s4 = reinterpret_cast<const float4*>(a.s);

	// accumulators: vertical candidates from the border row k = -1 (or the Local zero)
	float acc[NR][TP];
	#pragma unroll
	for (int i = 0; i < NR; i++) {
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			if (BORD) {
				const float top = LOC == VX_GLOBAL ? -a.wt[v + 1] : 0.0f;   // V[-1, v]  (solver.h:3108-3117)
				acc[i][v] = top - a.ws[u0 + i + 1];
			} else {
				acc[i][v] = 0.0f;
			}
		}
	}

	// vertical candidates from the finished rows k = 0 .. u0-1: rows [0, rt) from tensor memory, rows [rt, u0) from
	// shared memory.  u0 and rt are multiples of U = 4, so both ranges hold an even number of row PAIRS; the loops are
	// double-buffered by hand: the loads of the next pair are in flight while the 8 * TP candidates of this pair are
	// evaluated (with two warps per scheduler the load latency -- ~52 cycles for tcgen05.ld -- was exposed in every
	// iteration: 14-25 % long-scoreboard stalls in these loops, profiles/r2_wsb_f2).
	const int rt = a.rt;
	const int k_tmem = u0 < rt ? u0 : rt;
	if (PIPE) {
		float hA0[TP], hA1[TP], hB0[TP], hB1[TP];
		if (k_tmem > 0) {
			tmem_ld_pair<TP>(tmem, 0, hA0, hA1);
		}
		#pragma unroll 1
		for (int k = 0; k < k_tmem; k += 4) {
			tmem_wait_ld();                                   // pair A = rows k, k+1
			tmem_ld_pair<TP>(tmem, k + 2, hB0, hB1);          // in flight during A
			wsb_history_pair<TP, NR, F2>(a, u0 - k, acc, hA0, hA1);
			tmem_wait_ld();                                   // pair B = rows k+2, k+3
			if (k + 4 < k_tmem) {
				tmem_ld_pair<TP>(tmem, k + 4, hA0, hA1);      // in flight during B
			}
			wsb_history_pair<TP, NR, F2>(a, u0 - k - 2, acc, hB0, hB1);
		}
		if (rt < u0) {
			smem_ld_pair<TP>(hist, 0, hA0, hA1);
		}
		#pragma unroll 1
		for (int k = rt; k < u0; k += 4) {
			smem_ld_pair<TP>(hist, k - rt + 2, hB0, hB1);
			wsb_history_pair<TP, NR, F2>(a, u0 - k, acc, hA0, hA1);
			if (k + 4 < u0) {
				smem_ld_pair<TP>(hist, k - rt + 4, hA0, hA1);
			}
			wsb_history_pair<TP, NR, F2>(a, u0 - k - 2, acc, hB0, hB1);
		}
	} else {
		#pragma unroll 1
		for (int k = 0; k < k_tmem; k += 2) {
			float h0[TP], h1[TP];
			tmem_ld_pair<TP>(tmem, k, h0, h1);
			tmem_wait_ld();
			wsb_history_pair<TP, NR, F2>(a, u0 - k, acc, h0, h1);
		}
		#pragma unroll 1
		for (int k = rt; k < u0; k += 2) {
			float h0[TP], h1[TP];
			smem_ld_pair<TP>(hist, k - rt, h0, h1);
			wsb_history_pair<TP, NR, F2>(a, u0 - k, acc, h0, h1);
		}
	}

	// token ids: idq_next (ids of the next block, loaded one block ago) is consumed by the row-wise
	// S refills below; the ids of the block after that are requested now
	const int4 idq_use = idq_next;
	idq_next = __ldg(quads);
	quads += 32;
	uint32_t posq_next = 0;
	if (TAG) {
		posq_next = __ldg(pquads);
		pquads += 32;
	}

	// the rows of the block, in order
	#pragma unroll
	for (int i = 0; i < NR; i++) {
		const int u = u0 + i;
		// V[u, -1] and V[u-1, -1]
		const float left = BORD ? (LOC == VX_GLOBAL ? -a.ws[u + 1] : 0.0f) : 0.0f;
		float diag = BORD ? (LOC == VX_GLOBAL ? (u == 0 ? -a.wt[0] : -a.ws[u]) : 0.0f) : 0.0f;
		if (TAG) {
			vx_apply_tag<TP>(a.tag, (int8_t)((posq >> (8 * i)) & 0xffu), sim[i]);
		}
		if (F2) {
			// rows of this block above row i: pairs of rows and pairs of columns per FADD2; an odd row is left
			// pending and folds with the diagonal candidate below
			float pend[TP];
			#pragma unroll
			for (int j = 0; j + 1 < i; j += 2) {
				#pragma unroll
				for (int v = 0; v < TP; v += 2) {
					const float2 p = add2(acc[j][v], acc[j][v + 1], a.nws2[i - j]);
					const float2 q = add2(acc[j + 1][v], acc[j + 1][v + 1], a.nws2[i - j - 1]);
					acc[i][v] = max3(acc[i][v], p.x, q.x);
					acc[i][v + 1] = max3(acc[i][v + 1], p.y, q.y);
				}
			}
			if (i & 1) {
				#pragma unroll
				for (int v = 0; v < TP; v += 2) {
					const float2 p = add2(acc[i - 1][v], acc[i - 1][v + 1], a.nws2[1]);
					pend[v] = p.x;
					pend[v + 1] = p.y;
				}
			}
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				Fold f(acc[i][v]);
				if (i & 1) {
					f.push(pend[v]);
				}
				f.push(diag + sim[i][v]);
				diag = i == 0 ? prev[v] : acc[i - 1][v];
				if (BORD) {
					f.push(left - a.wt[v + 1]);
				}
				#pragma unroll
				for (int k = 0; k + 1 < v; k += 2) {
					const float2 p = add2(acc[i][k], acc[i][k + 1], a.nwt2[v - k]);   // cells k, k+1 to the left
					f.push2(p.x, p.y);
				}
				if (v & 1) {
					f.push(acc[i][v - 1] - a.wt[1]);
				}
				acc[i][v] = f.done();
			}
		} else {
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				Fold f(acc[i][v]);
				#pragma unroll
				for (int j = 0; j < i; j++) {
					f.push(acc[j][v] - a.ws[i - j]);          // rows of this block
				}
				f.push(diag + sim[i][v]);
				diag = i == 0 ? prev[v] : acc[i - 1][v];
				if (BORD) {
					f.push(left - a.wt[v + 1]);
				}
				#pragma unroll
				for (int k = 0; k < v; k++) {
					f.push(acc[i][k] - a.wt[v - k]);          // cells to the left
				}
				acc[i][v] = f.done();
			}
		}
		// this row of S has been consumed: fetch the same row of the next block
		load_sim_row<TP>(s4, quad_get(idq_use, i), sim[i]);

		if (LOC == VX_LOCAL) {
			constexpr int NF = TP - 3;   // columns [0, NF) always belong to the query (TP - 4 < t <= TP)
			constexpr int NP = NF / 2 * 2;
			#pragma unroll
			for (int v = 0; v < NP; v += 2) {
				best = max3(best, acc[i][v], acc[i][v + 1]);
			}
			#pragma unroll
			for (int v = NP; v < TP; v++) {
				if (v < NF || v < t) {
					best = fmaxf(best, acc[i][v]);
				}
			}
		} else {
			// V[u, t-1] (TP - 4 < t <= TP); the selects are opaque to the compiler, which would otherwise turn
			// the chain into a dynamically indexed -- i.e. local-memory -- copy of the accumulators
			float lc = acc[i][TP - 4];
			#pragma unroll
			for (int v = TP - 3; v < TP; v++) {
				asm("{ .reg .pred p; setp.eq.s32 p, %2, %3; selp.f32 %0, %1, %0, p; }"
					: "+f"(lc) : "f"(acc[i][v]), "r"(v), "r"(t - 1));
			}
			last_col = lc;
			if (LOC == VX_SEMIGLOBAL) {
				best = fmaxf(best, last_col);
			}
		}
	}
	#pragma unroll
	for (int i = NR; i < U; i++) {
		load_sim_row<TP>(s4, quad_get(idq_use, i), sim[i]);
	}

	// publish the rows
	if (u0 < rt) {
		#pragma unroll
		for (int i = 0; i < NR; i++) {
			#pragma unroll
			for (int q = 0; q < Q; q++) {
				tmem_st4(tmem + (uint32_t)((u0 + i) * TP + 4 * q), &acc[i][4 * q]);
			}
		}
		tmem_wait_st();
	} else {
		#pragma unroll
		for (int i = 0; i < NR; i++) {
			#pragma unroll
			for (int q = 0; q < Q; q++) {
				hist[((size_t)(u0 - rt + i) * Q + q) * 32] =
					make_float4(acc[i][4 * q + 0], acc[i][4 * q + 1], acc[i][4 * q + 2], acc[i][4 * q + 3]);
			}
		}
	}
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		prev[v] = acc[NR - 1][v];
	}
	if (TAG) {
		posq = posq_next;
	}
}

template<int TP, int LOC, bool BORD, bool F2, bool PIPE, bool TAG>
__global__ void __launch_bounds__(256, 1) wsb_score_kernel(const __grid_constant__ WsbArgs a) {
	extern __shared__ float4 wsb_smem[];
	__shared__ uint32_t tmem_base_sh;
	constexpr int Q = TP / 4;
	constexpr int U = VX_WSB_U;
	const int lane = threadIdx.x & 31;
	const int warp = threadIdx.x >> 5;
	const int hist_row0 = warp < a.n_long_warps ? warp * a.cap_long : a.n_long_warps * a.cap_long + (warp - a.n_long_warps) * a.cap;
	float4 *hist = wsb_smem + (size_t)hist_row0 * Q * 32 + lane;

	// let the next length class of this query start filling SMs as soon as CTAs of this launch retire
	// (programmatic dependent launch; there is no data dependency between the launches of one query)
	asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

	// all 512 TMEM columns of the SM (one CTA per SM); warp w owns lanes 32*(w%4).. and a column range
	if (warp == 0) {
		const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base_sh);
		const uint32_t ncols = 512;
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(dst), "r"(ncols) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
	const uint32_t tmem_base = tmem_base_sh;
	const uint32_t tmem = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(a.tmem_cols * (warp >> 2));

	const int t = a.t;
	const float4 *__restrict__ s4 = reinterpret_cast<const float4*>(a.s);

	for (;;) {
		// reserve a chunk of consecutive tiles; their token quads form one contiguous stream
		// Warps [0, n_long_warps) try the long queue first (stateless: one more atomic per reservation once it is empty;
		// a flag kept in a register for the whole kernel is what the 249-register variant has no room for).
		unsigned int got = 0;
		if (lane == 0) {
			bool lng = false;
			if ((threadIdx.x >> 5) < a.n_long_warps) {
				got = atomicAdd(a.counter_long, 1u);
				lng = got < a.n_long_tiles;
			}
			got = lng ? got | 0x80000000u : atomicAdd(a.counter, 1u);
		}
		got = __shfl_sync(0xffffffffu, got, 0);
		// guided schedule: large reservations first (the pipeline prologue is paid once per reservation), single
		// tiles at the end (the shortest windows: the warps finish within one short tile of each other)
		int tile, tile_stop;   // relative to tiles[0]: fewer than 2^26 tiles
		if (got & 0x80000000u) {
			tile = a.long_begin + (int)(got & 0x7fffffffu);
			tile_stop = tile + 1;
		} else {
			if (got < a.n_big) {
				tile = a.tile_begin + (int)got * a.chunk;
				tile_stop = tile + a.chunk;
			} else {
				tile = a.tile_begin + (int)a.n_big * a.chunk + (int)(got - a.n_big);
				tile_stop = tile + 1;
			}
			if (tile >= a.tile_end) {
				break;
			}
			tile_stop = tile_stop < a.tile_end ? tile_stop : a.tile_end;
		}
		VxTile tl = a.tiles[tile];
		const int4 *__restrict__ quads = reinterpret_cast<const int4*>(a.tok) + (size_t)tl.base128 * 32 + lane;

		// software pipeline over the stream: ids two blocks ahead, S rows one block ahead
		float sim[U][TP];
		int4 idq_next;
		{
			const int4 idq = __ldg(quads);
			idq_next = __ldg(quads + 32);
			#pragma unroll
			for (int i = 0; i < U; i++) {
				load_sim_row<TP>(s4, quad_get(idq, i), sim[i]);
			}
		}
		quads += 64;
		// POS bytes: this block's quad now, the next block's quad is requested inside wsb_block
		uint32_t posq = 0;
		const uint32_t *__restrict__ pquads = nullptr;
		if (TAG) {
			pquads = reinterpret_cast<const uint32_t*>(a.pos) + (size_t)tl.base128 * 32 + lane;
			posq = __ldg(pquads);
			pquads += 32;
		}

		for (; tile < tile_stop; tile++) {
			const VxTile tl_next = a.tiles[tile + 1];   // d_tiles has one padding entry
			const int len = tl.len;

			float prev[TP];
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				prev[v] = BORD && LOC == VX_GLOBAL ? -a.wt[v + 1] : 0.0f;   // V[-1, v]
			}
			float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
			float last_col = 0.0f;

			int u0 = 0;
			#pragma unroll 1
			for (; u0 + U <= len; u0 += U) {
				wsb_block<TP, LOC, BORD, U, F2, PIPE, TAG>(a, tmem, hist, u0, t, prev, best, last_col, sim, idq_next, quads, posq, pquads);
			}
			if (u0 < len) {
				switch (len - u0) {
					case 1:
						wsb_block<TP, LOC, BORD, 1, F2, PIPE, TAG>(a, tmem, hist, u0, t, prev, best, last_col, sim, idq_next, quads, posq, pquads);
						break;
					case 2:
						wsb_block<TP, LOC, BORD, 2, F2, PIPE, TAG>(a, tmem, hist, u0, t, prev, best, last_col, sim, idq_next, quads, posq, pquads);
						break;
					default:
						wsb_block<TP, LOC, BORD, 3, F2, PIPE, TAG>(a, tmem, hist, u0, t, prev, best, last_col, sim, idq_next, quads, posq, pquads);
						break;
				}
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
			const int64_t slot = (int64_t)tile * 32 + lane;
			const float boost = a.boost ? a.boost[slot] : 1.0f;
			a.score[slot] = lane < tl.n_valid ? (raw / a.total) * boost : -CUDART_INF_F;
			tl = tl_next;
			__syncwarp();
		}
	}

	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	if (warp == 0) {
		const uint32_t ncols = 512;
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem_base), "r"(ncols) : "memory");
	}
}

// ---------------------------------------------------------------------------
// host side: launch of one group of length classes (instantiated per query row width in vx_wsb_tp.cu)
// ---------------------------------------------------------------------------

// History placement for windows of up to `cap` rows: rows [0, rt) in tensor memory, the rest in shared
// memory.  Eight warps per SM (two per scheduler) when the rest fits; each pair of warps w, w+4 shares a
// TMEM lane quarter and gets 256 columns.  Long classes fall back to <= 4 warps with 512 columns each.
inline void wsb_plan(int tp, int cap, int *warps, int *rt, int *tmem_cols, int *rows_smem) {
	const size_t row_bytes = (size_t)tp * 32 * sizeof(float);
	const size_t smem_max = 227 * 1024 - 1024;
	*warps = 0;
	for (int w = 8; w >= 1; w--) {
		*tmem_cols = w > 4 ? 256 : 512;
		*rt = *tmem_cols / tp / VX_WSB_U * VX_WSB_U;
		*rows_smem = cap > *rt ? cap - *rt : 0;
		if ((size_t)*rows_smem * row_bytes * w <= smem_max) {
			*warps = w;
			return;
		}
	}
}

// long queue of a launch (see WsbArgs): n_warps warps per CTA sized for windows of up to `cap` rows take tiles [begin, end) first
struct WsbLong {
	int n_warps, cap;
	int64_t begin, end;
	unsigned int *counter;
};

template<int TP, int LOC, bool BORD, bool F2, bool PIPE, bool TAG>
static int launch_class(vx_corpus *c, WsbArgs &a, int cap, int sms, bool overlap, const WsbLong &lq) {
	// History placement: rows [0, rt) in tensor memory, the rest in shared memory.  Eight warps per
	// SM (two per scheduler) when the rest fits; each pair of warps w, w+4 shares a TMEM lane quarter
	// and gets 256 columns.  Long classes fall back to <= 4 warps with 512 columns each.
	int warps = 0, rt = 0, tmem_cols = 256, rows_smem = 0;
	wsb_plan(TP, cap, &warps, &rt, &tmem_cols, &rows_smem);
	const size_t row_bytes = (size_t)TP * 32 * sizeof(float);
	if (warps < 1) {
		return -1;
	}
	a.rt = rt;
	a.tmem_cols = tmem_cols;
	a.cap = rows_smem;
	const bool has_long = lq.n_warps > 0 && lq.end > lq.begin && warps == 8;
	const int rows_long = has_long ? lq.cap - rt : 0;
	a.n_long_warps = has_long ? lq.n_warps : 0;
	a.cap_long = rows_long;
	a.long_begin = has_long ? (int32_t)lq.begin : 0;
	a.n_long_tiles = has_long ? (uint32_t)(lq.end - lq.begin) : 0u;
	a.counter_long = lq.counter;
	const int64_t n_tiles = a.tile_end - a.tile_begin;
	// tiles per reservation: large enough to amortise the pipeline prologue, small enough that the
	// last chunks balance (>= 32 chunks per warp when the class is big)
	const int64_t lanes = (int64_t)sms * warps;
	const int64_t n_tail = n_tiles < 4 * lanes ? n_tiles : 4 * lanes;       // about four single tiles per warp at the end
	int64_t chunk = (n_tiles - n_tail) / (4 * lanes);                         // >= four large reservations per warp
	chunk = chunk < 1 ? 1 : chunk > 8 ? 8 : chunk;
	const int64_t n_big = (n_tiles - n_tail) / chunk;
	a.chunk = (int32_t)chunk;
	a.n_big = (uint32_t)n_big;
	const int64_t n_chunks = n_big + (n_tiles - n_big * chunk);
	const int64_t want_blocks = (n_chunks + warps - 1) / warps;
	int64_t need_blocks = want_blocks;
	if (has_long) {
		const int64_t long_blocks = (lq.end - lq.begin + lq.n_warps - 1) / lq.n_warps;
		need_blocks = long_blocks > need_blocks ? long_blocks : need_blocks;
	}
	const int blocks = (int)(need_blocks < sms ? need_blocks : sms);
	const size_t smem = ((size_t)rows_smem * (warps - a.n_long_warps) + (size_t)rows_long * a.n_long_warps) * row_bytes;
	static bool attr_set = false;
	if (!attr_set) {
		cudaFuncSetAttribute(wsb_score_kernel<TP, LOC, BORD, F2, PIPE, TAG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024 - 1024));
		attr_set = true;
	}
	// Every launch after the first of a query is a programmatic dependent launch: the kernels of one query do not
	// depend on each other (disjoint tiles and score slots, own chunk counter), so the CTAs of the next length
	// class move onto SMs as the CTAs of this one retire instead of waiting for its slowest CTA.
	cudaLaunchConfig_t cfg;
	memset(&cfg, 0, sizeof(cfg));
	cfg.gridDim = dim3((unsigned)blocks);
	cfg.blockDim = dim3((unsigned)(warps * 32));
	cfg.dynamicSmemBytes = smem;
	cfg.stream = c->stream;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = attr;
	cfg.numAttrs = overlap ? 1 : 0;
	if (cudaLaunchKernelEx(&cfg, wsb_score_kernel<TP, LOC, BORD, F2, PIPE, TAG>, a) != cudaSuccess) {
		return -1;
	}
	return 1;
}

template<int TP, bool F2, bool PIPE, bool TAG>
static int launch_class_f(vx_corpus *c, WsbArgs &a, int cap, int sms, int loc, bool bord, bool overlap, const WsbLong &lq) {
	if (loc == VX_LOCAL) {
		return bord ? launch_class<TP, VX_LOCAL, true, F2, PIPE, TAG>(c, a, cap, sms, overlap, lq) : launch_class<TP, VX_LOCAL, false, F2, PIPE, TAG>(c, a, cap, sms, overlap, lq);
	}
	if (loc == VX_GLOBAL) {
		return launch_class<TP, VX_GLOBAL, true, F2, PIPE, TAG>(c, a, cap, sms, overlap, lq);
	}
	return launch_class<TP, VX_SEMIGLOBAL, true, F2, PIPE, TAG>(c, a, cap, sms, overlap, lq);
}

template<int TP>
int vx_wsb_launch_class_t(vx_corpus *c, WsbArgs &a, int cap, int sms, int loc, bool bord, bool overlap, const WsbLong &lq) {
	// A/B switches for measurements (read once): VX_WSB_NO_F2 = scalar adds instead of FADD2 (and the round-1 loop
	// structure), VX_WSB_NO_PIPE = packed adds without the double-buffered history loops
	static const bool f2 = getenv("VX_WSB_NO_F2") == nullptr;
	static const bool pipe = getenv("VX_WSB_NO_PIPE") == nullptr;
	if (a.tag.on) {
		return launch_class_f<TP, true, false, true>(c, a, cap, sms, loc, bord, overlap, lq);   // tag weighting: packed adds, single-buffered
	}
	if (!f2) {
		return launch_class_f<TP, false, false, false>(c, a, cap, sms, loc, bord, overlap, lq);
	}
	// the second row-pair buffer costs 2 * TP registers: 249 at TP = 12 (fits 8 warps per SM exactly), spills at TP = 16
	if constexpr (TP <= 12) {
		if (pipe) {
			return launch_class_f<TP, true, true, false>(c, a, cap, sms, loc, bord, overlap, lq);
		}
	}
	return launch_class_f<TP, true, false, false>(c, a, cap, sms, loc, bord, overlap, lq);
}

