// Contextual similarity on the 5th-generation tensor cores.
//
// S_doc[e, 0..TP) = clip01(X[e, :] . Qhat[j, :]) for every token element e of the tiled corpus: the GEMM of
// ContextualEmbeddingSimilarityMatrixFactory (vectorian/core/cpp/metric/contextual.cpp:26-63) with the
// clip of ContextualEmbeddingSlice::similarity (cpp/slice/contextual.h:65-67) fused into the epilogue.
//
// The GEMM is [n_elem x d] . [d x 16]: 4*d bytes in and 4*TP bytes out per token, i.e. HBM-bound by a wide
// margin, but the FP32 SIMT version (ctx_sim_kernel) spends d*TP FMAs per token and tops out near half of
// the HBM rate.  Here the products run on tcgen05 with FP32 accuracy kept by the 3xTF32 split
//   x = x_hi + x_lo,  q = q_hi + q_lo,  x.q ~= x_hi.q_hi + x_lo.q_hi + x_hi.q_lo      (|dropped| <= 2^-22 |x||q|)
// with hi = x rounded to TF32 and lo = x - hi (exact in FP32).
//
// One persistent CTA per SM (17 warps), warp-specialised; a tile is 128 tokens, a sub-tile 32 floats of K:
//   warp  12    TMA producer: cp.async.bulk.tensor (SWIZZLE_128B box of 128 rows x 32 floats = 16 KB) of raw X
//               into a ring of up to 12 stages.  ~190 KB in flight per SM is what it takes to stream HBM at
//               full rate from one CTA per SM (scripts/ubench/readbw.cu: 64 KB in flight = 3.9 TB/s,
//               128 KB = 6.4, 256 KB = 7.2), far more than registers can hold
//   warps 0-7   converters: thread = row; LDS.128 of the row's 128 bytes (conflict-free thanks to the swizzle),
//               split into hi / lo in registers, tcgen05.st (32x32b.x32) into a 4-stage A ring in TMEM.
//               The stream never goes back to shared memory, which keeps smem traffic at 2 x 4*d bytes per token
//   warps 13-16 MMA issuers, one per 8-float k-step: tcgen05.mma.kind::tf32 with A from TMEM (M=128, K=8) and
//               Qhat from shared memory as one stacked operand [q_hi; q_lo] (N=32) resp. q_hi (N=16);
//               tcgen05.commit onto mbarriers frees the A stage / publishes the accumulators
//   warps 8-11  epilogue: tcgen05.ld of the accumulators (one TMEM lane quarter per warp, two ping-pong sets),
//               FP32 sum of the partial accumulators, clip to [0, 1], TP floats per row to S_doc
// Measured (5M tokens, d=128): 0.42-0.45 ms = 6.4 TB/s = 98% of the measured HBM copy rate; the FP32 SIMT
// kernel needs 0.73 ms (T<=8) / 1.12 ms (T<=16).  Max error vs float64 1e-7 (SIMT FP32: 2.5e-7).

#include "vx_internal.cuh"
#include <cuda.h>
#include <math_constants.h>
#include <vector>

namespace {

constexpr int CM_ROWS = 128;                 // tokens per tile (UMMA M)
constexpr int CM_SUB_BYTES = CM_ROWS * 128;  // one sub-tile: 128 rows x 32 floats
constexpr int CM_WARP_EPI = 8;               // warps 0-7 convert, 8-11 epilogue, 12 TMA, 13-16 MMA, 17 DP (fused)
constexpr int CM_WARP_TMA = 12;
constexpr int CM_WARP_MMA = 13;
constexpr int CM_WARP_DP = 17;               // fused kernels: the DP warp
constexpr int CM_THREADS = 18 * 32;
constexpr int CM_MAX_D = 768;
constexpr int CM_MAX_RAW = 12;
constexpr int CM_BCHUNK_BYTES = 32 * 128;    // Qhat per 32-float chunk of K: rows 0-15 = hi, rows 16-31 = lo, 128 B each

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
	return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
	asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
	asm volatile(
		"{\n\t"
		".reg .pred p;\n\t"
		"WAIT_%=:\n\t"
		"mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
		"@p bra DONE_%=;\n\t"
		"bra WAIT_%=;\n\t"
		"DONE_%=:\n\t"
		"}" :: "r"(bar), "r"(parity) : "memory");
}

__device__ __forceinline__ void umma_commit(uint32_t bar) {
	asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
	asm volatile(
		"{\n\t"
		".reg .pred p;\n\t"
		"setp.ne.b32 p, %4, 0;\n\t"
		"tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
		"}" :: "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
	asm volatile(
		"{\n\t"
		".reg .pred p;\n\t"
		"setp.ne.b32 p, %4, 0;\n\t"
		"tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t"
		"}" :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
	asm volatile(
		"tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
		"%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
		:: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
		   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
		   "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
		   "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
	asm volatile(
		"tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
		"%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
		: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
		  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
		  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
		  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
		: "r"(taddr));
}

// K-major SWIZZLE_128B shared-memory descriptor (cute/arch/mma_sm100_desc.hpp: start >> 4 in [0,14),
// LBO [16,30) = 1 (unused for swizzled K-major), SBO [32,46) = 1024 B >> 4 between 8-row groups,
// version [46,48) = 1, layout [61,64) = 2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
	return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// instruction descriptor: D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both K-major, N = 16, M = 128
constexpr uint32_t CM_IDESC_N16 = (1u << 4) | (2u << 7) | (2u << 10) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t CM_IDESC_N32 = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);

__device__ __forceinline__ void split_tf32(float x, float &hi, float &lo) {
	// round to nearest at 10 mantissa bits; inf / nan do not occur in normalised embeddings
	hi = __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
	const float r = x - hi;
	lo = __uint_as_float((__float_as_uint(r) + 0x1000u) & 0xffffe000u);
}

struct CtxMmaArgs {
	const float *qhat;   // [t][d]
	const int32_t *qid;  // static embeddings: vocabulary id of query token j (identity fix-up), else nullptr
	float *sdoc;         // [n_elem][tp]
	int64_t n_elem;
	int32_t d, t, tp, n_tiles;
	int32_t n_raw;       // stages of the raw-X ring in shared memory
	// fused linear-gap DP (FUSED kernels only): the corpus' window tiles and the scoring parameters
	const VxTile *tiles;
	int32_t n_vxtiles, n_quads;
	const float *boost;
	float *score;
	float gs, gt, total;
	int32_t locality, write_sdoc;
};

// TMEM map (columns): [0, 128) accumulator set 0, [128, 256) set 1 (each: 4 x {16 hi.hi, 16 cross}),
// [256, 512) ring of 4 A stages, each 32 columns x_hi + 32 columns x_lo of one 32-float chunk of K
constexpr int CM_ACC_COLS = 128;
constexpr int CM_A_COL0 = 256;
constexpr int CM_NA = 4;

// The DP warp of the fused kernel (see ctx_mma_kernel): windows tiles [i_lo, i_hi), S tiles from the two shared-memory
// buffers at st0 in quad order.
template<int TP, int LOC_>
__device__ __forceinline__ void ctx_dp_warp(const CtxMmaArgs &a, int i_lo, int i_hi, const float4 *st0,
	uint32_t bar_s_full, uint32_t bar_s_empty, int lane) {

	const float gs = a.gs, gt = a.gt;
	const int t = a.t;
	const int LOC = LOC_ < 0 ? a.locality : LOC_;   // LOC_ < 0: decided at run time
	uint32_t lt = 0;
	for (int i = i_lo; i < i_hi; i++) {
		const VxTile tl = a.tiles[i];
		const int len = tl.len;
		float h[TP];   // h[v] = V[u-1, v]
		#pragma unroll
		for (int v = 0; v < TP; v++) {
			h[v] = LOC == VX_GLOBAL ? -((float)(v + 1) * gt) : 0.0f;   // solver.h:2804-2806
		}
		float best = LOC == VX_LOCAL ? 0.0f : -CUDART_INF_F;
		float last_col = 0.0f;
		for (int u0 = 0; u0 < len; u0 += 4, lt++) {
			const uint32_t sb = lt & 1;
			mbar_wait(bar_s_full + 8 * sb, (lt >> 1) & 1);
			// The buffer of the PREVIOUS quad is released here, behind this wait: its loads have long been consumed
			// by the arithmetic (an arrive right behind the loads would not be ordered after them, see the
			// converters).  The epilogue needs that buffer one whole GEMM tile later, so nothing waits for it.
			if (lt > 0) {
				__syncwarp();
				if (lane == 0) {
					mbar_arrive(bar_s_empty + 8 * (sb ^ 1));
				}
			}
			const float4 *const st = st0 + sb * (CM_ROWS * TP / 4);
			#pragma unroll
			for (int k = 0; k < 4; k++) {
				const int u = u0 + k;
				if (u < len) {
					float cur[TP];
					#pragma unroll
					for (int jq = 0; jq < TP / 4; jq++) {
						const float4 x = st[(k * (TP / 4) + jq) * 32 + lane];
						cur[4 * jq + 0] = x.x;
						cur[4 * jq + 1] = x.y;
						cur[4 * jq + 2] = x.z;
						cur[4 * jq + 3] = x.w;
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
						if (LOC != VX_LOCAL && v == t - 1) {
							last_col = x;
						}
					}
					if (LOC == VX_SEMIGLOBAL) {
						best = fmaxf(best, last_col);
					}
				}
			}
		}
		float rawv;
		if (LOC == VX_LOCAL) {
			rawv = best;
		} else if (LOC == VX_GLOBAL) {
			rawv = last_col;
		} else {
			#pragma unroll
			for (int v = 0; v < TP; v++) {
				if (v < t) {
					best = fmaxf(best, h[v]);
				}
			}
			rawv = best;
		}
		const int64_t slot = (int64_t)i * 32 + lane;
		const float boost = a.boost ? a.boost[slot] : 1.0f;
		a.score[slot] = lane < tl.n_valid ? (rawv / a.total) * boost : -CUDART_INF_F;
	}
}

// FUSED: the similarity tiles feed the linear-gap DP (NW / SW / semiglobal) inside the same kernel.  A 128-token GEMM
// tile is exactly one token quad of 32 windows (tokens 4q .. 4q+3 of every lane of a window tile, vx_internal.cuh), so
// the epilogue warps drop the clipped tile into shared memory as [k][column quad][lane] float4 and ONE extra warp runs
// the DP of the 32 windows over it (lane = window, the same arithmetic as score_linear_kernel), quad after quad.  Each
// CTA owns a contiguous range of window tiles balanced by token quads, so S_doc never has to be read back.
template<int TP, bool FUSED>
__global__ void __launch_bounds__(CM_THREADS, 1) ctx_mma_kernel(const __grid_constant__ CUtensorMap tmap, const CtxMmaArgs a) {
	extern __shared__ uint8_t smem_raw[];
	const uint32_t raw = smem_u32(smem_raw);
	const uint32_t base = (raw + 1023u) & ~1023u;          // SWIZZLE_128B tiles want 1024-byte alignment
	uint8_t *const sm = smem_raw + (base - raw);

	const int n_chunk = (a.d + 31) / 32;                   // 32-float chunks of K = sub-tiles per tile (the last may be partial)
	const uint32_t n_raw = a.n_raw;
	const uint32_t off_b = n_raw * CM_SUB_BYTES;           // Qhat operand: n_chunk x 4 KB
	const uint32_t off_bar = off_b + n_chunk * CM_BCHUNK_BYTES;
	// barriers: raw_full[16], raw_empty[16], a_full[4], a_empty[4], acc_full[2], acc_empty[2], TMEM base
	const uint32_t bar_raw_full = base + off_bar, bar_raw_empty = bar_raw_full + 8 * 16;
	const uint32_t bar_a_full = bar_raw_empty + 8 * 16, bar_a_empty = bar_a_full + 8 * CM_NA;
	const uint32_t bar_acc_full = bar_a_empty + 8 * CM_NA, bar_acc_empty = bar_acc_full + 16;
	uint32_t *const tmem_slot = reinterpret_cast<uint32_t*>(sm + off_bar + 8 * (32 + 2 * CM_NA + 4));
	// + 384: qid[16]; + 448: s_full[2], s_empty[2]; + 480: fused tile range; + 512: fused S tiles (2 x 128 x TP floats)
	const uint32_t bar_s_full = base + off_bar + 448, bar_s_empty = bar_s_full + 16;
	const uint32_t off_stile = off_bar + 512;

	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

	if (threadIdx.x == 0) {
		for (uint32_t s = 0; s < n_raw; s++) {
			mbar_init(bar_raw_full + 8 * s, 1);            // the TMA producer's arrive.expect_tx
			mbar_init(bar_raw_empty + 8 * s, 4);           // one arrive per converter warp of the sub-tile
		}
		for (int s = 0; s < CM_NA; s++) {
			mbar_init(bar_a_full + 8 * s, 4);
			mbar_init(bar_a_empty + 8 * s, 4);             // tcgen05.commit of each MMA warp
		}
		for (int b = 0; b < 2; b++) {
			mbar_init(bar_acc_full + 8 * b, 4);            // tcgen05.commit of each MMA warp
			mbar_init(bar_acc_empty + 8 * b, 4);           // one arrive per epilogue warp
		}
		for (int b = 0; b < 2; b++) {
			mbar_init(bar_s_full + 8 * b, 4);              // fused: S tile written by the four epilogue warps
			mbar_init(bar_s_empty + 8 * b, 1);             // fused: consumed by the DP warp
		}
		if (FUSED) {
			// window tiles [i_lo, i_hi) of this CTA: boundaries at equal shares of the token quads
			int32_t *rng = reinterpret_cast<int32_t*>(sm + off_bar + 480);
			for (int e = 0; e < 2; e++) {
				const int64_t target = (int64_t)a.n_quads * (blockIdx.x + e) / gridDim.x;
				int lo = 0, hi = a.n_vxtiles;      // first tile with base128 >= target
				while (lo < hi) {
					const int mid = (lo + hi) >> 1;
					if ((int64_t)a.tiles[mid].base128 < target) {
						lo = mid + 1;
					} else {
						hi = mid;
					}
				}
				rng[e] = blockIdx.x + e == gridDim.x ? a.n_vxtiles : lo;
			}
			rng[2] = rng[0] < a.n_vxtiles ? (int32_t)a.tiles[rng[0]].base128 : a.n_quads;
			rng[3] = rng[1] < a.n_vxtiles ? (int32_t)a.tiles[rng[1]].base128 : a.n_quads;
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if (warp == CM_WARP_MMA) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(tmem_slot)) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	// Qhat -> one K-major SWIZZLE_128B operand of 32 rows per chunk: rows 0-15 hi, rows 16-31 lo (rows >= t zero)
	const int d_pad = 32 * n_chunk;                        // X columns >= d read as zero (TMA out-of-bounds fill)
	for (int i = threadIdx.x; i < 16 * d_pad; i += blockDim.x) {
		const int n = i / d_pad, k = i - n * d_pad;
		float hi = 0.0f, lo = 0.0f;
		if (n < a.t && k < a.d) {
			split_tf32(__ldg(a.qhat + (size_t)n * a.d + k), hi, lo);
		}
		const int chunk = k >> 5, kc = (k >> 2) & 7, w = k & 3;
		const uint32_t o = off_b + chunk * CM_BCHUNK_BYTES + n * 128 + ((kc ^ (n & 7)) << 4) + 4 * w;
		*reinterpret_cast<float*>(sm + o) = hi;
		*reinterpret_cast<float*>(sm + o + 16 * 128) = lo;
	}
	if (threadIdx.x < 16) {
		reinterpret_cast<int32_t*>(sm + off_bar + 384)[threadIdx.x] =
			a.qid && (int)threadIdx.x < a.t ? __ldg(a.qid + threadIdx.x) : -1;
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
	const uint32_t tmem = *tmem_slot;
	int tile_first, tile_step, my_tiles;
	if (FUSED) {
		const int32_t *rng = reinterpret_cast<const int32_t*>(sm + off_bar + 480);
		tile_first = rng[2];
		tile_step = 1;
		my_tiles = rng[3] - rng[2];
	} else {
		tile_first = blockIdx.x;
		tile_step = gridDim.x;
		my_tiles = (int)blockIdx.x < a.n_tiles ? (a.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
	}
	const uint32_t n_sub = (uint32_t)my_tiles * n_chunk;   // sub-tiles this CTA processes, in order j = lt * n_chunk + c
	if (warp == CM_WARP_TMA) {
		// ---- TMA producer: raw FP32 sub-tiles of X, up to n_raw x 16 KB in flight ----
		if (lane == 0) {
			uint32_t s = 0, ph = 0;
			int c = 0, tile = tile_first;
			for (uint32_t j = 0; j < n_sub; j++) {
				mbar_wait(bar_raw_empty + 8 * s, ph ^ 1);
				asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
					:: "r"(bar_raw_full + 8 * s), "r"((uint32_t)CM_SUB_BYTES) : "memory");
				asm volatile(
					"cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
					:: "r"(base + s * CM_SUB_BYTES), "l"(&tmap), "r"(32 * c), "r"(tile * CM_ROWS),
					   "r"(bar_raw_full + 8 * s) : "memory");
				if (++s == n_raw) {
					s = 0;
					ph ^= 1;
				}
				if (++c == n_chunk) {
					c = 0;
					tile += tile_step;
				}
			}
		}
	} else if (warp >= CM_WARP_MMA && warp < CM_WARP_MMA + 4) {
		// ---- four MMA issuers, one per 8-float step k8 of a sub-tile: A (x_hi, x_lo) from TMEM, B (Qhat hi | lo)
		// from shared memory.  One warp issuing all eight MMAs of a sub-tile was the bottleneck (~130 mostly
		// uniform-datapath instructions per sub-tile); the k8 steps use disjoint accumulators, so they split cleanly.
		const int k8 = warp - CM_WARP_MMA;
		// x_hi . [q_hi; q_lo] leaves hi.hi in columns [0, 16) and hi.lo in [16, 32) of accumulator k8; x_lo . q_hi is
		// added to the latter.  The tensor core truncates on every accumulation, so the large hi.hi sums are spread
		// over four accumulators (partial sums stay small) and added in FP32 by the epilogue.
		const uint64_t db0 = umma_desc(base + off_b + 32 * k8);
		int c = 0;
		uint32_t lt = 0, ts = 0, pa = 0;
		for (uint32_t j = 0; j < n_sub; j++) {
			const uint32_t b = lt & 1;
			if (c == 0) {
				mbar_wait(bar_acc_empty + 8 * b, ((lt >> 1) & 1) ^ 1);
			}
			mbar_wait(bar_a_full + 8 * ts, pa);
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			if (lane == 0) {
				const uint32_t d_t = tmem + CM_ACC_COLS * b + 32 * k8;
				const uint32_t a_t = tmem + CM_A_COL0 + 64 * ts + 8 * k8;
				const uint64_t db = db0 + (uint64_t)(c * (CM_BCHUNK_BYTES >> 4));
				umma_tf32_ts(d_t, a_t, db, CM_IDESC_N32, c != 0);
				umma_tf32_ts(d_t + 16, a_t + 32, db, CM_IDESC_N16, 1);
				umma_commit(bar_a_empty + 8 * ts);
				if (c == n_chunk - 1) {
					umma_commit(bar_acc_full + 8 * b);
				}
			}
			__syncwarp();
			if (++c == n_chunk) {
				c = 0;
				lt++;
			}
			if (++ts == CM_NA) {
				ts = 0;
				pa ^= 1;
			}
		}
	} else if (warp < CM_WARP_EPI) {
		// ---- converters: raw X (smem) -> x_hi, x_lo (TMEM); thread = row, two warps per TMEM lane quarter ----
		const int q = warp & 3, h = warp >> 2;
		const int row = 32 * q + lane;
		for (uint32_t j = h; j < n_sub; j += 2) {
			const uint32_t s = j % n_raw, ph = (j / n_raw) & 1;
			mbar_wait(bar_raw_full + 8 * s, ph);
			const uint8_t *const rp = sm + s * CM_SUB_BYTES + row * 128;
			float4 v[8];
			#pragma unroll
			for (int kc = 0; kc < 8; kc++) {
				v[kc] = *reinterpret_cast<const float4*>(rp + ((kc ^ (row & 7)) << 4));
			}
			uint32_t hi[32], lo[32];
			#pragma unroll
			for (int kc = 0; kc < 8; kc++) {
				const float x[4] = {v[kc].x, v[kc].y, v[kc].z, v[kc].w};
				#pragma unroll
				for (int w = 0; w < 4; w++) {
					float fh, fl;
					split_tf32(x[w], fh, fl);
					hi[4 * kc + w] = __float_as_uint(fh);
					lo[4 * kc + w] = __float_as_uint(fl);
				}
			}
			const uint32_t ts = j % CM_NA, pa = (j / CM_NA) & 1;
			mbar_wait(bar_a_empty + 8 * ts, pa ^ 1);
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + CM_A_COL0 + 64 * ts;
			tmem_st32(taddr, hi);
			tmem_st32(taddr + 32, lo);
			// Release the raw stage only now: the tcgen05.st have read hi / lo, so the eight LDS.128 above have
			// delivered their data.  An arrive issued right after the loads is NOT ordered behind them (SASS: LDS.128 x 8,
			// SYNCS.ARRIVE with nothing in between) and the TMA refill of the stage could land before the loads had read
			// it: seen as a rare wrong S row under back-pressure (fused kernel, L2-resident X; scripts/stress_ctx.py).
			if (lane == 0) {
				mbar_arrive(bar_raw_empty + 8 * s);
			}
			asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
			asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
			__syncwarp();
			if (lane == 0) {
				mbar_arrive(bar_a_full + 8 * ts);
			}
		}
	} else if (warp < CM_WARP_TMA) {
		// ---- epilogue: accumulators -> clip01 -> S_doc (and, fused, the shared-memory tile of the DP warp) ----
		const int q = warp & 3;                 // TMEM lane quarter this warp may read
		const int32_t *const qid = reinterpret_cast<const int32_t*>(sm + off_bar + 384);   // filled before the CTA barrier
		for (uint32_t lt = 0; lt < (uint32_t)my_tiles; lt++) {
			const int tile = tile_first + (int)lt * tile_step;
			const uint32_t b = lt & 1, u = lt >> 1;
			mbar_wait(bar_acc_full + 8 * b, u & 1);
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + CM_ACC_COLS * b;
			float hh[16], cr[16];
			#pragma unroll
			for (int g = 0; g < 4; g++) {
				uint32_t r[32];
				tmem_ld32(taddr + 32 * g, r);
				asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
				#pragma unroll
				for (int j = 0; j < 16; j++) {
					hh[j] = g == 0 ? __uint_as_float(r[j]) : hh[j] + __uint_as_float(r[j]);
					cr[j] = g == 0 ? __uint_as_float(r[16 + j]) : cr[j] + __uint_as_float(r[16 + j]);
				}
			}
			asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
			__syncwarp();
			if (lane == 0) {
				mbar_arrive(bar_acc_empty + 8 * b);     // after tcgen05.wait::ld: the accumulators are in registers
			}
			const int row = 32 * q + lane;
			const int64_t e = (int64_t)tile * CM_ROWS + row;
			float4 o4[TP / 4];
			#pragma unroll
			for (int jq = 0; jq < TP / 4; jq++) {
				float o_[4];
				#pragma unroll
				for (int w = 0; w < 4; w++) {
					// identity fix-up (cpp/metric/static.cpp:58-67), then clip (cpp/metric/metric.h:28-30)
					const float x = (int64_t)qid[4 * jq + w] == e ? 1.0f : hh[4 * jq + w] + cr[4 * jq + w];
					o_[w] = fminf(fmaxf(x, 0.0f), 1.0f);
				}
				o4[jq] = make_float4(o_[0], o_[1], o_[2], o_[3]);
			}
			if (e < a.n_elem && (!FUSED || a.write_sdoc)) {
				float4 *const dst = reinterpret_cast<float4*>(a.sdoc + e * TP);
				#pragma unroll
				for (int jq = 0; jq < TP / 4; jq++) {
					dst[jq] = o4[jq];
				}
			}
			if (FUSED) {
				// row = 4 * window lane + token k of the quad -> [k][jq][window lane]: the DP warp reads it conflict-free
				const uint32_t sb = lt & 1;
				mbar_wait(bar_s_empty + 8 * sb, ((lt >> 1) & 1) ^ 1);
				float4 *const st = reinterpret_cast<float4*>(sm + off_stile) + sb * (CM_ROWS * TP / 4);
				#pragma unroll
				for (int jq = 0; jq < TP / 4; jq++) {
					st[((row & 3) * (TP / 4) + jq) * 32 + (row >> 2)] = o4[jq];
				}
				__syncwarp();
				if (lane == 0) {
					mbar_arrive(bar_s_full + 8 * sb);
				}
			}
		}
	} else if (FUSED) {
		// ---- DP warp: linear-gap alignment of the 32 windows of every window tile, lane = window
		// (LinearGapCostSolver::solve, pyalign/solver.h:2817-2847; the arithmetic of score_linear_kernel, vx_align.cu).
		// For T > 8 the locality is a compile-time constant of the loop (one instantiation per locality, selected once per
		// CTA): the per-cell selects of a run-time locality cost a third of the warp's instructions, and this single warp
		// is what bounds the kernel at T = 16 (C4-like case: 0.55 -> 0.48 ms).
		const int32_t *rng = reinterpret_cast<const int32_t*>(sm + off_bar + 480);
		const float4 *const st0 = reinterpret_cast<const float4*>(sm + off_stile);
		if constexpr (TP >= 12) {
			if (a.locality == VX_LOCAL) {
				ctx_dp_warp<TP, VX_LOCAL>(a, rng[0], rng[1], st0, bar_s_full, bar_s_empty, lane);
			} else if (a.locality == VX_GLOBAL) {
				ctx_dp_warp<TP, VX_GLOBAL>(a, rng[0], rng[1], st0, bar_s_full, bar_s_empty, lane);
			} else {
				ctx_dp_warp<TP, VX_SEMIGLOBAL>(a, rng[0], rng[1], st0, bar_s_full, bar_s_empty, lane);
			}
		} else {
			// T <= 8: the warp is far from bounding the kernel, and the three copies measured slower (0.50 vs 0.44 ms
			// on the C4-like case: the kernel is sensitive to its code size), so the locality stays a run-time value
			ctx_dp_warp<TP, -1>(a, rng[0], rng[1], st0, bar_s_full, bar_s_empty, lane);
		}
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	if (warp == CM_WARP_MMA) {
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
	}
}

} // namespace

typedef CUresult (*vx_tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
	const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
	CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct CtxFusedParams {
	const VxTile *tiles;
	int32_t n_vxtiles, n_quads;
	const float *boost;
	float *score;
	float gs, gt, total;
	int32_t locality, write_sdoc;
};

// returns 1 when the tensor-core kernel was launched, 0 when the shape is outside what it handles
// (the caller then uses the SIMT kernel), -1 on an unsupported tp or a failed tensor-map encode.
// fused != nullptr: the linear-gap DP of every window runs inside the same kernel (ctx_mma_kernel<TP, true>).
static int launch_ctx_mma(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, const int32_t *d_qid,
	int32_t t, int32_t tp, float *d_sdoc, const CtxFusedParams *fused, cudaStream_t st) {

	if (d % 4 != 0 || d > CM_MAX_D || n_elem <= 0 || (reinterpret_cast<uintptr_t>(d_x) & 15) != 0) {
		return 0;
	}
	static int sms = 0;
	static vx_tmap_encode_fn encode = nullptr;
	if (sms == 0) {
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
		void *fn = nullptr;
		cudaDriverEntryPointQueryResult qres;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
			qres != cudaDriverEntryPointSuccess) {
			fn = nullptr;
		}
		encode = reinterpret_cast<vx_tmap_encode_fn>(fn);
	}
	if (!encode) {
		return 0;
	}
	// X as a 2-d tensor [n_elem][d], box = 32 floats x 128 rows, 128-byte swizzle; rows past n_elem read as zero
	CUtensorMap tmap;
	const cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)n_elem};
	const cuuint64_t gstride[1] = {(cuuint64_t)d * sizeof(float)};
	const cuuint32_t box[2] = {32, CM_ROWS};
	const cuuint32_t estr[2] = {1, 1};
	if (encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(d_x), gdim, gstride, box, estr,
		CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
		CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
		return -1;
	}
	CtxMmaArgs a;
	memset(&a, 0, sizeof(a));
	a.qhat = d_qhat;
	a.qid = d_qid;
	a.sdoc = d_sdoc;
	a.n_elem = n_elem;
	a.d = d;
	a.t = t;
	a.tp = tp;
	a.n_tiles = (int32_t)((n_elem + CM_ROWS - 1) / CM_ROWS);
	int grid = a.n_tiles < sms ? a.n_tiles : sms;
	size_t fixed = 1024 + (size_t)((d + 31) / 32) * CM_BCHUNK_BYTES + 512;
	if (fused) {
		a.tiles = fused->tiles;
		a.n_vxtiles = fused->n_vxtiles;
		a.n_quads = fused->n_quads;
		a.boost = fused->boost;
		a.score = fused->score;
		a.gs = fused->gs;
		a.gt = fused->gt;
		a.total = fused->total;
		a.locality = fused->locality;
		a.write_sdoc = fused->write_sdoc;
		grid = fused->n_vxtiles < sms ? fused->n_vxtiles : sms;
		fixed += 2 * (size_t)CM_ROWS * tp * sizeof(float);   // the two S tiles of the DP warp
	}
	int n_raw = (int)((227 * 1024 - fixed) / CM_SUB_BYTES);
	n_raw = n_raw > CM_MAX_RAW ? CM_MAX_RAW : n_raw;
	// even: sub-tile j goes to converter group j & 1 and to stage j % n_raw, so with an even ring a stage always belongs
	// to the same group and every warp sees every phase of its barriers (with an odd ring a warp would skip every other
	// phase and a parity wait could be satisfied by the phase before the one it means)
	n_raw &= ~1;
	if (n_raw < 2 || grid < 1) {
		return 0;
	}
	a.n_raw = n_raw;
	const size_t smem = fixed + (size_t)n_raw * CM_SUB_BYTES;
	#define VX_CM_CASE(TP) case TP: \
		if (fused) { \
			static size_t smem_set = 0; \
			if (smem > smem_set) { \
				cudaFuncSetAttribute(ctx_mma_kernel<TP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
				smem_set = smem; \
			} \
			ctx_mma_kernel<TP, true><<<grid, CM_THREADS, smem, st>>>(tmap, a); \
		} else { \
			static size_t smem_set = 0; \
			if (smem > smem_set) { \
				cudaFuncSetAttribute(ctx_mma_kernel<TP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
				smem_set = smem; \
			} \
			ctx_mma_kernel<TP, false><<<grid, CM_THREADS, smem, st>>>(tmap, a); \
		} \
		break;
	switch (tp) {
		VX_CM_CASE(4)
		VX_CM_CASE(8)
		VX_CM_CASE(12)
		VX_CM_CASE(16)
		default:
			return -1;
	}
	#undef VX_CM_CASE
	return 1;
}

int vx_launch_ctx_mma(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, const int32_t *d_qid,
	int32_t t, int32_t tp, float *d_sdoc, cudaStream_t st) {
	return launch_ctx_mma(d_x, n_elem, d, d_qhat, d_qid, t, tp, d_sdoc, nullptr, st);
}

// Contextual query with a linear gap cost: similarity GEMM and the DP of every window in one kernel.  S_doc is still
// written (the traceback of the k winners reads it) but never read back by the scoring pass.  Returns like above.
int vx_launch_ctx_fused(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, float *d_sdoc) {
	if (ap.gap_kind != VX_GAP_LINEAR || c->n_tiles <= 0 || c->n_tiles > 0x7fffffff / 64) {
		return 0;
	}
	CtxFusedParams f;
	f.tiles = c->d_tiles;
	f.n_vxtiles = (int32_t)c->n_tiles;
	f.n_quads = (int32_t)((c->n_tok_tiled - VX_TOK_SLACK) / 128);
	f.boost = c->d_boost;
	f.score = c->d_score;
	f.gs = ap.lin_s;
	f.gt = ap.lin_t;
	f.total = qp.total;
	f.locality = ap.locality;
	f.write_sdoc = 1;
	return launch_ctx_mma(c->d_ctx, c->n_tok_tiled, c->ctx_dim, s->d_qhat, nullptr, s->t, s->tp, d_sdoc, &f, c->stream);
}

// test hook (include/vectorian_b200.h): both contextual similarity kernels on host data
extern "C" int vx_debug_ctx_sim(const float *x, int64_t n_elem, int32_t d, const float *qhat, int32_t t, int32_t tp,
	int32_t kernel, float *sdoc_out, float *ms) {

	if (!x || !qhat || !sdoc_out || n_elem <= 0 || d <= 0 || t <= 0 || t > tp) {
		VX_FAIL(VX_ERR_INVALID, "vx_debug_ctx_sim: bad argument");
	}
	float *d_x = nullptr, *d_q = nullptr, *d_s = nullptr;
	cudaEvent_t e0, e1;
	VX_CUDA(cudaMalloc(&d_x, sizeof(float) * (size_t)n_elem * d));
	VX_CUDA(cudaMalloc(&d_q, sizeof(float) * (size_t)t * d));
	// 4 KB guard zones before and after S_doc: an out-of-bounds store of either kernel fails the hook
	const size_t guard = 1024;
	float *d_s_alloc = nullptr;
	VX_CUDA(cudaMalloc(&d_s_alloc, sizeof(float) * ((size_t)n_elem * tp + 2 * guard)));
	VX_CUDA(cudaMemset(d_s_alloc, 0x5a, sizeof(float) * ((size_t)n_elem * tp + 2 * guard)));
	d_s = d_s_alloc + guard;
	VX_CUDA(cudaMemcpy(d_x, x, sizeof(float) * (size_t)n_elem * d, cudaMemcpyHostToDevice));
	VX_CUDA(cudaMemcpy(d_q, qhat, sizeof(float) * (size_t)t * d, cudaMemcpyHostToDevice));
	VX_CUDA(cudaMemset(d_s, 0xff, sizeof(float) * (size_t)n_elem * tp));
	VX_CUDA(cudaEventCreate(&e0));
	VX_CUDA(cudaEventCreate(&e1));
	int rc = 0;
	for (int rep = 0; rep < 2; rep++) {   // the second launch is the timed one
		VX_CUDA(cudaEventRecord(e0, 0));
		if (kernel == 1) {
			rc = vx_launch_ctx_mma(d_x, n_elem, d, d_q, nullptr, t, tp, d_s, 0);
		} else {
			vx_ctx_force_simt = 1;
			rc = vx_launch_ctx_sim(d_x, n_elem, d, d_q, t, tp, d_s, 0);
			vx_ctx_force_simt = 0;
		}
		VX_CUDA(cudaEventRecord(e1, 0));
		if (rc != 1) {
			break;
		}
	}
	cudaError_t err = cudaDeviceSynchronize();
	if (rc == 1 && err == cudaSuccess) {
		if (ms) {
			cudaEventElapsedTime(ms, e0, e1);
		}
		err = cudaMemcpy(sdoc_out, d_s, sizeof(float) * (size_t)n_elem * tp, cudaMemcpyDeviceToHost);
	}
	bool guard_ok = true;
	if (err == cudaSuccess) {
		std::vector<uint32_t> g(2 * guard);
		err = cudaMemcpy(g.data(), d_s_alloc, sizeof(float) * guard, cudaMemcpyDeviceToHost);
		if (err == cudaSuccess) {
			err = cudaMemcpy(g.data() + guard, d_s + (size_t)n_elem * tp, sizeof(float) * guard, cudaMemcpyDeviceToHost);
		}
		for (uint32_t x : g) {
			guard_ok = guard_ok && x == 0x5a5a5a5au;
		}
	}
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(d_x);
	cudaFree(d_q);
	cudaFree(d_s_alloc);
	if (err != cudaSuccess) {
		VX_FAIL(VX_ERR_CUDA, std::string("vx_debug_ctx_sim: ") + cudaGetErrorString(err));
	}
	if (!guard_ok) {
		VX_FAIL(VX_ERR_CUDA, "vx_debug_ctx_sim: a kernel wrote outside S_doc");
	}
	if (rc != 1) {
		VX_FAIL(VX_ERR_INVALID, "vx_debug_ctx_sim: shape not handled by this kernel");
	}
	return VX_OK;
}
