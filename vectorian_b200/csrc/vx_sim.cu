// Similarity-matrix kernels.
//
// Replaces CosineSim.compute (vectorian/sim/vector.py:66-78), Vectors.normalized / magnitudes
// (vectorian/embedding/vectors.py:71-86) and build_static_similarity_matrix
// (vectorian/core/cpp/metric/static.cpp:9-78): S = clip01(Ehat . Qhat^T) with S[id(q_j), j] = 1.
//
// For one query the contraction is V x d x T with T <= 16: 2*T flop per 4 bytes of E, far below the
// tensor-core ridge, so this is an HBM-streaming FP32 kernel (E is read exactly once).
#include "vx_internal.cuh"

__device__ __forceinline__ float warp_sum(float x) {
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		x += __shfl_xor_sync(0xffffffffu, x, o);
	}
	return x;
}

// One warp per row: out = row / |row|, 0/0 -> 0 (nan_to_num).
__global__ void normalize_rows_kernel(const float *__restrict__ in, float *__restrict__ out,
	int64_t rows, int32_t d) {

	const int64_t row = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
	const int lane = threadIdx.x & 31;
	if (row >= rows) {
		return;
	}
	const float *src = in + row * d;
	float acc = 0.0f;
	for (int k = lane; k < d; k += 32) {
		const float x = src[k];
		acc = fmaf(x, x, acc);
	}
	acc = warp_sum(acc);
	float mag = sqrtf(acc);
	if (isnan(mag)) {
		mag = 0.0f;
	}
	float *dst = out + row * d;
	for (int k = lane; k < d; k += 32) {
		float x = src[k] / mag;
		if (isnan(x)) {
			x = 0.0f;
		}
		dst[k] = x;
	}
}

int vx_launch_normalize(const float *d_in, float *d_out, int64_t rows, int32_t d, cudaStream_t st) {
	if (rows == 0) {
		return 0;
	}
	const int warps = 8;
	normalize_rows_kernel<<<(unsigned)((rows + warps - 1) / warps), warps * 32, 0, st>>>(d_in, d_out, rows, d);
	return 1;
}

// 8 lanes per row, 2 rows per lane group: lane `sub` of a group takes the dimensions k = 8m + sub, so
// the group reads 32 contiguous bytes of each row per step and -- with Qhat staged k-major in shared
// memory ([d][TPS], TPS padded so that the 8 lanes hit 8 different bank quads) -- one LDS.128 feeds
// 4 FFMA of each of the 2 rows.  3 shuffle steps reduce a row.  E is read exactly once: HBM-bound.
template<int TP>
__global__ void __launch_bounds__(256) sim_static_g8_kernel(const float *__restrict__ ehat, int64_t v, int32_t d,
	const float *__restrict__ qhat, const int32_t *__restrict__ qid, int32_t t,
	float *__restrict__ s) {

	constexpr int TPS = TP % 8 == 0 ? TP + 4 : TP;
	extern __shared__ float q_sh[];   // [d][TPS], columns >= t are zero
	for (int i = threadIdx.x; i < d * TP; i += blockDim.x) {
		const int k = i / TP, j = i - k * TP;
		q_sh[k * TPS + j] = j < t ? qhat[(size_t)j * d + k] : 0.0f;
	}
	__syncthreads();

	const int lane = threadIdx.x & 31;
	const int sub = lane & 7;
	const int grp = lane >> 3;
	const int64_t warps_total = (int64_t)gridDim.x * (blockDim.x >> 5);
	const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);

	for (int64_t row0 = warp * 8; row0 < v; row0 += warps_total * 8) {
		const int64_t ra = row0 + grp * 2, rb = ra + 1;
		const float *__restrict__ pa = ehat + (ra < v ? ra : v - 1) * d;
		const float *__restrict__ pb = ehat + (rb < v ? rb : v - 1) * d;
		float acc_a[TP], acc_b[TP];
		#pragma unroll
		for (int j = 0; j < TP; j++) {
			acc_a[j] = 0.0f;
			acc_b[j] = 0.0f;
		}
		#pragma unroll 4
		for (int k = sub; k < d; k += 8) {
			const float xa = __ldg(pa + k);
			const float xb = __ldg(pb + k);
			const float4 *__restrict__ q4 = reinterpret_cast<const float4*>(q_sh + k * TPS);
			#pragma unroll
			for (int jq = 0; jq < TP / 4; jq++) {
				const float4 qv = q4[jq];
				acc_a[4 * jq + 0] = fmaf(xa, qv.x, acc_a[4 * jq + 0]);
				acc_a[4 * jq + 1] = fmaf(xa, qv.y, acc_a[4 * jq + 1]);
				acc_a[4 * jq + 2] = fmaf(xa, qv.z, acc_a[4 * jq + 2]);
				acc_a[4 * jq + 3] = fmaf(xa, qv.w, acc_a[4 * jq + 3]);
				acc_b[4 * jq + 0] = fmaf(xb, qv.x, acc_b[4 * jq + 0]);
				acc_b[4 * jq + 1] = fmaf(xb, qv.y, acc_b[4 * jq + 1]);
				acc_b[4 * jq + 2] = fmaf(xb, qv.z, acc_b[4 * jq + 2]);
				acc_b[4 * jq + 3] = fmaf(xb, qv.w, acc_b[4 * jq + 3]);
			}
		}
		#pragma unroll
		for (int j = 0; j < TP; j++) {
			#pragma unroll
			for (int o = 4; o > 0; o >>= 1) {
				acc_a[j] += __shfl_xor_sync(0xffffffffu, acc_a[j], o);
				acc_b[j] += __shfl_xor_sync(0xffffffffu, acc_b[j], o);
			}
		}
		if (sub == 0) {
			#pragma unroll
			for (int r = 0; r < 2; r++) {
				const int64_t row = ra + r;
				if (row < v) {
					float out[TP];
					#pragma unroll
					for (int j = 0; j < TP; j++) {
						float x = 0.0f;
						if (j < t) {
							x = r == 0 ? acc_a[j] : acc_b[j];
							if (qid[j] == row) {
								x = 1.0f;                     // identity fix-up, cpp/metric/static.cpp:58-67
							}
							x = fminf(fmaxf(x, 0.0f), 1.0f);  // clip, cpp/metric/metric.h:28-30
						}
						out[j] = x;
					}
					float4 *dst = reinterpret_cast<float4*>(s + row * TP);
					#pragma unroll
					for (int jq = 0; jq < TP / 4; jq++) {
						dst[jq] = make_float4(out[4 * jq], out[4 * jq + 1], out[4 * jq + 2], out[4 * jq + 3]);
					}
				}
			}
		}
	}
}

int vx_launch_sim_static(const float *d_ehat, int64_t v, int32_t d, const float *d_qhat,
	const int32_t *d_qid, int32_t t, int32_t tp, float *d_s, cudaStream_t st) {

	if (v == 0) {
		return 0;
	}
	// tensor-core path (vx_ctx_mma.cu): the same GEMM with the identity fix-up in the epilogue
	static const bool simt_env = getenv("VX_SIM_SIMT") != nullptr;   // A/B switch for measurements
	if (!simt_env && !vx_ctx_force_simt) {
		const int rc = vx_launch_ctx_mma(d_ehat, v, d, d_qhat, d_qid, t, tp, d_s, st);
		if (rc != 0) {
			return rc;
		}
	}
	const int warps = 8;
	{
		const size_t smem4 = sizeof(float) * (size_t)d * (tp % 8 == 0 ? tp + 4 : tp);
		int64_t blocks4 = (v + warps * 8 - 1) / (warps * 8);
		const int64_t cap4 = 148 * (smem4 > 48 * 1024 ? 2 : 6);
		if (blocks4 > cap4) {
			blocks4 = cap4;
		}
		// the attribute is raised once per instantiation and only ever grows (one process drives one device)
		#define VX_SIM4_CASE(TP) case TP: { \
			static size_t smem_set = 0; \
			if (smem4 > smem_set) { \
				cudaFuncSetAttribute(sim_static_g8_kernel<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4); \
				smem_set = smem4; \
			} } \
			sim_static_g8_kernel<TP><<<(unsigned)blocks4, warps * 32, smem4, st>>>(d_ehat, v, d, d_qhat, d_qid, t, d_s); \
			return 1;
		switch (tp) {
			VX_SIM4_CASE(4)
			VX_SIM4_CASE(8)
			VX_SIM4_CASE(12)
			VX_SIM4_CASE(16)
			default:
				return -1;
		}
		#undef VX_SIM4_CASE
	}
}


// ---------------------------------------------------------------------------
// Contextual similarity: S_doc[e, 0..TP) = clip01(X[e, :] . Qhat[j, :]) for every token element e of the
// tiled corpus (replaces the per-document GEMM of ContextualEmbeddingSimilarityMatrixFactory,
// vectorian/core/cpp/metric/contextual.cpp:26-63, and ContextualEmbeddingSlice::similarity,
// cpp/slice/contextual.h:65-67).  FP32 SIMT version: every thread owns two elements, streams their
// d floats with 16-byte loads (the rows are contiguous, each fetched line is consumed by the same
// thread over the next 8 loads) and takes Qhat k-major from shared memory as broadcast LDS.128.
// X is read exactly once: 4*d bytes in, 4*TP bytes out per token.
// ---------------------------------------------------------------------------

template<int TP>
__global__ void __launch_bounds__(128) ctx_sim_kernel(const float *__restrict__ x, int64_t n_elem, int32_t d,
	const float *__restrict__ qhat, int32_t t, float *__restrict__ sdoc) {

	extern __shared__ float q_sh[];   // [d][TP], columns >= t are zero
	for (int i = threadIdx.x; i < d * TP; i += blockDim.x) {
		const int k = i / TP, j = i - k * TP;
		q_sh[i] = j < t ? qhat[(size_t)j * d + k] : 0.0f;
	}
	__syncthreads();
	const float4 *__restrict__ q4 = reinterpret_cast<const float4*>(q_sh);
	const int d4 = d >> 2;

	for (int64_t base = (int64_t)blockIdx.x * 256; base < n_elem; base += (int64_t)gridDim.x * 256) {
		const int64_t ea = base + threadIdx.x, eb = ea + 128;
		const float4 *__restrict__ pa = reinterpret_cast<const float4*>(x + (ea < n_elem ? ea : n_elem - 1) * d);
		const float4 *__restrict__ pb = reinterpret_cast<const float4*>(x + (eb < n_elem ? eb : n_elem - 1) * d);
		float acc_a[TP], acc_b[TP];
		#pragma unroll
		for (int j = 0; j < TP; j++) {
			acc_a[j] = 0.0f;
			acc_b[j] = 0.0f;
		}
		#pragma unroll 4
		for (int i = 0; i < d4; i++) {
			const float4 xa = __ldg(pa + i);
			const float4 xb = __ldg(pb + i);
			const float xa_[4] = {xa.x, xa.y, xa.z, xa.w};
			const float xb_[4] = {xb.x, xb.y, xb.z, xb.w};
			#pragma unroll
			for (int kk = 0; kk < 4; kk++) {
				#pragma unroll
				for (int jq = 0; jq < TP / 4; jq++) {
					const float4 qv = q4[(4 * i + kk) * (TP / 4) + jq];
					acc_a[4 * jq + 0] = fmaf(xa_[kk], qv.x, acc_a[4 * jq + 0]);
					acc_a[4 * jq + 1] = fmaf(xa_[kk], qv.y, acc_a[4 * jq + 1]);
					acc_a[4 * jq + 2] = fmaf(xa_[kk], qv.z, acc_a[4 * jq + 2]);
					acc_a[4 * jq + 3] = fmaf(xa_[kk], qv.w, acc_a[4 * jq + 3]);
					acc_b[4 * jq + 0] = fmaf(xb_[kk], qv.x, acc_b[4 * jq + 0]);
					acc_b[4 * jq + 1] = fmaf(xb_[kk], qv.y, acc_b[4 * jq + 1]);
					acc_b[4 * jq + 2] = fmaf(xb_[kk], qv.z, acc_b[4 * jq + 2]);
					acc_b[4 * jq + 3] = fmaf(xb_[kk], qv.w, acc_b[4 * jq + 3]);
				}
			}
		}
		#pragma unroll
		for (int r = 0; r < 2; r++) {
			const int64_t e = r == 0 ? ea : eb;
			if (e < n_elem) {
				float4 *dst = reinterpret_cast<float4*>(sdoc + e * TP);
				#pragma unroll
				for (int jq = 0; jq < TP / 4; jq++) {
					float o[4];
					#pragma unroll
					for (int c = 0; c < 4; c++) {
						const float v = r == 0 ? acc_a[4 * jq + c] : acc_b[4 * jq + c];
						o[c] = 4 * jq + c < t ? fminf(fmaxf(v, 0.0f), 1.0f) : 0.0f;   // clip, contextual.cpp:52-56
					}
					dst[jq] = make_float4(o[0], o[1], o[2], o[3]);
				}
			}
		}
	}
}

// generic d (not a multiple of 4): one thread per element, scalar loads
template<int TP>
__global__ void __launch_bounds__(128) ctx_sim_scalar_kernel(const float *__restrict__ x, int64_t n_elem, int32_t d,
	const float *__restrict__ qhat, int32_t t, float *__restrict__ sdoc) {

	const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (e >= n_elem) {
		return;
	}
	const float *px = x + e * d;
	#pragma unroll
	for (int j = 0; j < TP; j++) {
		float acc = 0.0f;
		if (j < t) {
			for (int k = 0; k < d; k++) {
				acc = fmaf(px[k], qhat[(size_t)j * d + k], acc);
			}
			acc = fminf(fmaxf(acc, 0.0f), 1.0f);
		}
		sdoc[e * TP + j] = acc;
	}
}

int vx_ctx_force_simt = 0;   // set by the vx_debug_ctx_sim test hook

int vx_launch_ctx_sim(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, int32_t t, int32_t tp,
	float *d_sdoc, cudaStream_t st) {

	if (n_elem == 0) {
		return 0;
	}
	static const bool simt_env = getenv("VX_CTX_SIMT") != nullptr;   // A/B switch for measurements
	if (!simt_env && !vx_ctx_force_simt) {
		const int rc = vx_launch_ctx_mma(d_x, n_elem, d, d_qhat, nullptr, t, tp, d_sdoc, st);
		if (rc != 0) {
			return rc;
		}
	}
	const bool vec = d % 4 == 0 && (size_t)d * tp * sizeof(float) <= 200 * 1024;
	const size_t smem = sizeof(float) * (size_t)d * tp;
	int64_t blocks = (n_elem + 255) / 256;
	if (blocks > 148 * 8) {
		blocks = 148 * 8;
	}
	#define VX_CTX_CASE(TP) case TP: \
		if (vec) { \
			static size_t smem_set = 0; \
			if (smem > smem_set) { \
				cudaFuncSetAttribute(ctx_sim_kernel<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
				smem_set = smem; \
			} \
			ctx_sim_kernel<TP><<<(unsigned)blocks, 128, smem, st>>>(d_x, n_elem, d, d_qhat, t, d_sdoc); \
		} else { \
			ctx_sim_scalar_kernel<TP><<<(unsigned)((n_elem + 127) / 128), 128, 0, st>>>(d_x, n_elem, d, d_qhat, t, d_sdoc); \
		} \
		break;
	switch (tp) {
		VX_CTX_CASE(4)
		VX_CTX_CASE(8)
		VX_CTX_CASE(12)
		VX_CTX_CASE(16)
		default:
			return -1;
	}
	#undef VX_CTX_CASE
	return 1;
}

__global__ void iota_kernel(int32_t *out, int64_t n) {
	for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
		out[i] = (int32_t)i;
	}
}

int vx_launch_iota(int32_t *d_out, int64_t n, cudaStream_t st) {
	if (n == 0) {
		return 0;
	}
	iota_kernel<<<148 * 8, 256, 0, st>>>(d_out, n);
	return 1;
}
