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
// One persistent CTA per SM, warp-specialised:
//   warps 0-7   producers: coalesced LDG.128 of a [128 rows x 64 floats] slab of X, split into hi / lo in
//               registers and stored to shared memory in the K-major SWIZZLE_128B canonical layout
//               (what a TMA box load with CU_TENSOR_MAP_SWIZZLE_128B would have produced, but the split has
//               to pass through registers anyway, so the loads do as well)
//   warp  8     MMA issuer: 24 x tcgen05.mma.kind::tf32 (M=128, N=16, K=8) per slab into TMEM accumulators
//               (two ping-pong sets of five 16-column accumulators), tcgen05.commit onto mbarriers
//   warps 9-12  epilogue: tcgen05.ld of the 128 x 16 accumulator (one TMEM lane quarter per warp), clip to
//               [0, 1], TP floats per row to S_doc
// Shared-memory ring: 3 stages x 64 KB (hi + lo), Qhat hi / lo resident (d * 128 B).

#include "vx_internal.cuh"

namespace {

constexpr int CM_ROWS = 128;                 // tokens per tile (UMMA M)
constexpr int CM_KS = 64;                    // floats of K per stage: two SWIZZLE_128B sub-tiles
constexpr int CM_SUB_BYTES = CM_ROWS * 128;  // one sub-tile: 128 rows x 128 B
constexpr int CM_STAGE_BYTES = 4 * CM_SUB_BYTES;   // hi sub0, hi sub1, lo sub0, lo sub1
constexpr int CM_NSTAGE = 3;
constexpr int CM_PRODUCERS = 256;
constexpr int CM_THREADS = CM_PRODUCERS + 32 + 128;
constexpr int CM_MAX_D = 256;
constexpr int CM_ACC_COLS = 128;              // TMEM columns per accumulator set (5 x 16 used), two sets
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
	const float *x;      // [n_elem][d]
	const float *qhat;   // [t][d]
	float *sdoc;         // [n_elem][tp]
	int64_t n_elem;
	int32_t d, t, tp, n_tiles;
};

template<int TP>
__global__ void __launch_bounds__(CM_THREADS, 1) ctx_mma_kernel(const CtxMmaArgs a) {
	extern __shared__ uint8_t smem_raw[];
	const uint32_t raw = smem_u32(smem_raw);
	const uint32_t base = (raw + 1023u) & ~1023u;          // SWIZZLE_128B tiles want 1024-byte alignment
	uint8_t *const sm = smem_raw + (base - raw);

	const int n_chunk = a.d / 32;                          // 32-float chunks of K
	const int n_stage_k = a.d / CM_KS;                     // slabs per tile
	const uint32_t off_b_hi = CM_NSTAGE * CM_STAGE_BYTES;
	const uint32_t off_b_lo = off_b_hi + 16 * 128;
	const uint32_t off_bar = off_b_hi + n_chunk * CM_BCHUNK_BYTES;
	// barriers: full[3], empty[3], acc_full[2], acc_empty[2], then the TMEM base address
	const uint32_t bar_full = base + off_bar, bar_empty = bar_full + 8 * CM_NSTAGE;
	const uint32_t bar_acc_full = bar_empty + 8 * CM_NSTAGE, bar_acc_empty = bar_acc_full + 16;
	uint32_t *const tmem_slot = reinterpret_cast<uint32_t*>(sm + off_bar + 8 * (2 * CM_NSTAGE + 4));

	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

	if (threadIdx.x == 0) {
		for (int s = 0; s < CM_NSTAGE; s++) {
			mbar_init(bar_full + 8 * s, CM_PRODUCERS);
			mbar_init(bar_empty + 8 * s, 1);
		}
		for (int b = 0; b < 2; b++) {
			mbar_init(bar_acc_full + 8 * b, 1);
			mbar_init(bar_acc_empty + 8 * b, 128);
		}
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	if (warp == 8) {
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" :: "r"(smem_u32(tmem_slot)) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	// Qhat -> hi / lo tiles in the same swizzled K-major layout (rows >= t are zero)
	for (int i = threadIdx.x; i < 16 * a.d; i += CM_THREADS) {
		const int n = i / a.d, k = i - n * a.d;
		float hi = 0.0f, lo = 0.0f;
		if (n < a.t) {
			split_tf32(__ldg(a.qhat + (size_t)n * a.d + k), hi, lo);
		}
		const int chunk = k >> 5, kc = (k >> 2) & 7, w = k & 3;
		const uint32_t o = chunk * CM_BCHUNK_BYTES + n * 128 + ((kc ^ (n & 7)) << 4) + 4 * w;
		*reinterpret_cast<float*>(sm + off_b_hi + o) = hi;
		*reinterpret_cast<float*>(sm + off_b_lo + o) = lo;
	}
	asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
	const uint32_t tmem = *tmem_slot;
	if (warp < 8) {
		// ---- producers ----
		// software pipeline: the loads of slab it + 1 are in flight while slab it is split and stored
		auto load_slab = [&](int tile, int ks, float4 (&v)[8]) {
			const int64_t row0 = (int64_t)tile * CM_ROWS;
			#pragma unroll
			for (int i = 0; i < 8; i++) {
				const int f = i * CM_PRODUCERS + threadIdx.x;
				const int row = f >> 4, c16 = f & 15;
				int64_t e = row0 + row;
				e = e < a.n_elem ? e : a.n_elem - 1;
				v[i] = __ldg(reinterpret_cast<const float4*>(a.x + e * a.d + ks * CM_KS) + c16);
			}
		};
		auto store_slab = [&](uint32_t it, const float4 (&v)[8]) {
			const uint32_t s = it % CM_NSTAGE, ph = (it / CM_NSTAGE) & 1;
			mbar_wait(bar_empty + 8 * s, ph ^ 1);
			uint8_t *const st = sm + s * CM_STAGE_BYTES;
			#pragma unroll
			for (int i = 0; i < 8; i++) {
				const int f = i * CM_PRODUCERS + threadIdx.x;
				const int row = f >> 4, c16 = f & 15;
				const uint32_t o = (c16 >> 3) * CM_SUB_BYTES + row * 128 + (((c16 & 7) ^ (row & 7)) << 4);
				float4 hi, lo;
				split_tf32(v[i].x, hi.x, lo.x);
				split_tf32(v[i].y, hi.y, lo.y);
				split_tf32(v[i].z, hi.z, lo.z);
				split_tf32(v[i].w, hi.w, lo.w);
				*reinterpret_cast<float4*>(st + o) = hi;
				*reinterpret_cast<float4*>(st + 2 * CM_SUB_BYTES + o) = lo;
			}
			asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
			mbar_arrive(bar_full + 8 * s);
		};
		// slab sequence of this CTA: it -> (tile, ks); two slabs are in flight while a third is stored
		const int my_tiles = blockIdx.x < a.n_tiles ? (a.n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
		const uint32_t n_it = (uint32_t)my_tiles * n_stage_k;
		auto load_it = [&](uint32_t it, float4 (&v)[8]) {
			if (it < n_it) {
				const int lt = it / n_stage_k;
				load_slab(blockIdx.x + lt * gridDim.x, it - lt * n_stage_k, v);
			}
		};
		float4 v0[8], v1[8], v2[8];
		load_it(0, v0);
		load_it(1, v1);
		for (uint32_t it = 0; it < n_it; it += 3) {
			load_it(it + 2, v2);
			store_slab(it, v0);
			if (it + 1 >= n_it) {
				break;
			}
			load_it(it + 3, v0);
			store_slab(it + 1, v1);
			if (it + 2 >= n_it) {
				break;
			}
			load_it(it + 4, v1);
			store_slab(it + 2, v2);
		}
	} else if (warp == 8) {
		// ---- MMA issuer ----
		uint32_t it = 0, lt = 0;
		for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, lt++) {
			const uint32_t b = lt & 1, u = lt >> 1;
			mbar_wait(bar_acc_empty + 8 * b, (u & 1) ^ 1);
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			const uint32_t d_tmem = tmem + CM_ACC_COLS * b;
			for (int ks = 0; ks < n_stage_k; ks++, it++) {
				const uint32_t s = it % CM_NSTAGE, ph = (it / CM_NSTAGE) & 1;
				mbar_wait(bar_full + 8 * s, ph);
				asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
				if (lane == 0) {
					const uint32_t a_hi = base + s * CM_STAGE_BYTES, a_lo = a_hi + 2 * CM_SUB_BYTES;
					#pragma unroll
					for (int sub = 0; sub < 2; sub++) {
						const uint32_t b_hi = base + off_b_hi + (2 * ks + sub) * CM_BCHUNK_BYTES;
						#pragma unroll
						for (int k8 = 0; k8 < 4; k8++) {
							const uint32_t ao = sub * CM_SUB_BYTES + 32 * k8;
							const uint64_t dah = umma_desc(a_hi + ao), dal = umma_desc(a_lo + ao);
							const uint64_t db = umma_desc(b_hi + 32 * k8);
							// Qhat hi and lo are one 32-row operand: x_hi . [q_hi; q_lo] leaves hi.hi in columns
							// [0, 16) and hi.lo in [16, 32) of the accumulator; x_lo . q_hi is added to the latter.
							// The tensor core truncates on every accumulation, so the large hi.hi sums are kept in
							// four accumulators (one per k8: partial sums stay small) and added in FP32 afterwards.
							umma_tf32(d_tmem + 32 * k8, dah, db, CM_IDESC_N32, (ks | sub) != 0);
							umma_tf32(d_tmem + 32 * k8 + 16, dal, db, CM_IDESC_N16, 1);
						}
					}
					umma_commit(bar_empty + 8 * s);
					if (ks == n_stage_k - 1) {
						umma_commit(bar_acc_full + 8 * b);
					}
				}
				__syncwarp();
			}
		}
	} else {
		// ---- epilogue ----
		const int q = warp & 3;                 // TMEM lane quarter this warp may read
		uint32_t lt = 0;
		for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, lt++) {
			const uint32_t b = lt & 1, u = lt >> 1;
			mbar_wait(bar_acc_full + 8 * b, u & 1);
			asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
			const uint32_t taddr = tmem + ((uint32_t)(32 * q) << 16) + CM_ACC_COLS * b;
			float hh[16], cr[16];
			#pragma unroll
			for (int g = 0; g < 4; g++) {
				uint32_t r[32];
				asm volatile(
					"tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
					"%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
					: "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
					  "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
					  "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
					  "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
					: "r"(taddr + 32 * g));
				asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
				#pragma unroll
				for (int j = 0; j < 16; j++) {
					hh[j] = g == 0 ? __uint_as_float(r[j]) : hh[j] + __uint_as_float(r[j]);
					cr[j] = g == 0 ? __uint_as_float(r[16 + j]) : cr[j] + __uint_as_float(r[16 + j]);
				}
			}
			asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
			mbar_arrive(bar_acc_empty + 8 * b);
			const int64_t e = (int64_t)tile * CM_ROWS + 32 * q + lane;
			if (e < a.n_elem) {
				float4 *const dst = reinterpret_cast<float4*>(a.sdoc + e * TP);
				#pragma unroll
				for (int jq = 0; jq < TP / 4; jq++) {
					float o_[4];
					#pragma unroll
					for (int w = 0; w < 4; w++) {
						o_[w] = fminf(fmaxf(hh[4 * jq + w] + cr[4 * jq + w], 0.0f), 1.0f);
					}
					const float4 o = make_float4(o_[0], o_[1], o_[2], o_[3]);
					dst[jq] = o;
				}
			}
		}
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	if (warp == 8) {
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" :: "r"(tmem) : "memory");
	}
}

} // namespace

// returns 1 when the tensor-core kernel was launched, 0 when the shape is outside what it handles
// (the caller then uses the SIMT kernel), -1 on an unsupported tp
int vx_launch_ctx_mma(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, int32_t t, int32_t tp,
	float *d_sdoc, cudaStream_t st) {

	if (d % CM_KS != 0 || d > CM_MAX_D || n_elem < CM_ROWS) {
		return 0;
	}
	static int sms = 0;
	if (sms == 0) {
		int dev = 0;
		cudaGetDevice(&dev);
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
	}
	CtxMmaArgs a;
	a.x = d_x;
	a.qhat = d_qhat;
	a.sdoc = d_sdoc;
	a.n_elem = n_elem;
	a.d = d;
	a.t = t;
	a.tp = tp;
	a.n_tiles = (int32_t)((n_elem + CM_ROWS - 1) / CM_ROWS);
	const size_t smem = 1024 + (size_t)CM_NSTAGE * CM_STAGE_BYTES + (size_t)(d / 32) * CM_BCHUNK_BYTES + 128;
	const int grid = a.n_tiles < sms ? a.n_tiles : sms;
	#define VX_CM_CASE(TP) case TP: \
		cudaFuncSetAttribute(ctx_mma_kernel<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
		ctx_mma_kernel<TP><<<grid, CM_THREADS, smem, st>>>(a); \
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
	VX_CUDA(cudaMalloc(&d_s, sizeof(float) * (size_t)n_elem * tp));
	VX_CUDA(cudaMemcpy(d_x, x, sizeof(float) * (size_t)n_elem * d, cudaMemcpyHostToDevice));
	VX_CUDA(cudaMemcpy(d_q, qhat, sizeof(float) * (size_t)t * d, cudaMemcpyHostToDevice));
	VX_CUDA(cudaMemset(d_s, 0xff, sizeof(float) * (size_t)n_elem * tp));
	VX_CUDA(cudaEventCreate(&e0));
	VX_CUDA(cudaEventCreate(&e1));
	int rc = 0;
	for (int rep = 0; rep < 2; rep++) {   // the second launch is the timed one
		VX_CUDA(cudaEventRecord(e0, 0));
		if (kernel == 1) {
			rc = vx_launch_ctx_mma(d_x, n_elem, d, d_q, t, tp, d_s, 0);
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
	cudaEventDestroy(e0);
	cudaEventDestroy(e1);
	cudaFree(d_x);
	cudaFree(d_q);
	cudaFree(d_s);
	if (err != cudaSuccess) {
		VX_FAIL(VX_ERR_CUDA, std::string("vx_debug_ctx_sim: ") + cudaGetErrorString(err));
	}
	if (rc != 1) {
		VX_FAIL(VX_ERR_INVALID, "vx_debug_ctx_sim: shape not handled by this kernel");
	}
	return VX_OK;
}
