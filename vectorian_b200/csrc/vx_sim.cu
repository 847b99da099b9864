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

// S row per warp.  Qhat[t x d] is staged in shared memory once per block; each lane keeps TP partial
// dot products over its strided slice of the row, then the warp reduces them.
template<int TP>
__global__ void sim_static_kernel(const float *__restrict__ ehat, int64_t v, int32_t d,
	const float *__restrict__ qhat, const int32_t *__restrict__ qid, int32_t t,
	float *__restrict__ s) {

	extern __shared__ float q_sh[];   // [t x d]
	for (int i = threadIdx.x; i < t * d; i += blockDim.x) {
		q_sh[i] = qhat[i];
	}
	__syncthreads();

	const int lane = threadIdx.x & 31;
	const int warps_per_block = blockDim.x / 32;
	for (int64_t row = (int64_t)blockIdx.x * warps_per_block + threadIdx.x / 32; row < v;
		row += (int64_t)gridDim.x * warps_per_block) {

		const float *src = ehat + row * d;
		float acc[TP];
		#pragma unroll
		for (int j = 0; j < TP; j++) {
			acc[j] = 0.0f;
		}
		for (int k = lane; k < d; k += 32) {
			const float x = src[k];
			#pragma unroll
			for (int j = 0; j < TP; j++) {
				if (j < t) {
					acc[j] = fmaf(x, q_sh[j * d + k], acc[j]);
				}
			}
		}
		#pragma unroll
		for (int j = 0; j < TP; j++) {
			acc[j] = warp_sum(acc[j]);
		}
		if (lane == 0) {
			#pragma unroll
			for (int j = 0; j < TP; j++) {
				float x = 0.0f;
				if (j < t) {
					x = acc[j];
					if (qid[j] == row) {
						x = 1.0f;                     // identity fix-up, cpp/metric/static.cpp:58-67
					}
					x = fminf(fmaxf(x, 0.0f), 1.0f);  // clip, cpp/metric/metric.h:28-30
				}
				s[row * TP + j] = x;
			}
		}
	}
}

int vx_launch_sim_static(const float *d_ehat, int64_t v, int32_t d, const float *d_qhat,
	const int32_t *d_qid, int32_t t, int32_t tp, float *d_s, cudaStream_t st) {

	if (v == 0) {
		return 0;
	}
	const int warps = 8;
	const size_t smem = sizeof(float) * (size_t)t * d;
	int64_t blocks = (v + warps - 1) / warps;
	if (blocks > 148 * 16) {
		blocks = 148 * 16;
	}
	#define VX_SIM_CASE(TP) case TP: \
		cudaFuncSetAttribute(sim_static_kernel<TP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
		sim_static_kernel<TP><<<(unsigned)blocks, warps * 32, smem, st>>>(d_ehat, v, d, d_qhat, d_qid, t, d_s); \
		break;
	switch (tp) {
		VX_SIM_CASE(4)
		VX_SIM_CASE(8)
		VX_SIM_CASE(12)
		VX_SIM_CASE(16)
		default:
			return -1;
	}
	#undef VX_SIM_CASE
	return 1;
}
