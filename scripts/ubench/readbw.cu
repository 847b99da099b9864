// Read-bandwidth calibration for the contextual similarity kernels: how fast can X be streamed at all?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o readbw readbw.cu && ./readbw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

// (a) grid-stride LDG.128, UNR loads in flight per thread
template<int UNR>
__global__ void read_stride(const float4 *x, size_t n4, float *out) {
	float acc = 0.0f;
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	for (; i + (UNR - 1) * stride < n4; i += UNR * stride) {
		float4 v[UNR];
		#pragma unroll
		for (int u = 0; u < UNR; u++) v[u] = __ldg(x + i + u * stride);
		#pragma unroll
		for (int u = 0; u < UNR; u++) acc += v[u].x + v[u].y + v[u].z + v[u].w;
	}
	if (acc == 123.456f) out[0] = acc;
}

// (b) persistent CTAs, 64 KB tiles (128 rows x 512 B), 256 threads, half a row-slab (256 B per row) per step,
// two steps prefetched: the access pattern of ctx_mma_kernel's producers without any of the rest
__global__ void __launch_bounds__(256) read_tiles(const float4 *x, int n_tiles, float *out) {
	float acc = 0.0f;
	for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
		const float4 *base = x + (size_t)tile * 128 * 32;
		float4 v[16];
		#pragma unroll
		for (int ks = 0; ks < 2; ks++) {
			#pragma unroll
			for (int i = 0; i < 8; i++) {
				const int f = i * 256 + threadIdx.x;
				v[ks * 8 + i] = __ldg(base + (f >> 4) * 32 + ks * 16 + (f & 15));
			}
		}
		#pragma unroll
		for (int i = 0; i < 16; i++) acc += v[i].x + v[i].y + v[i].z + v[i].w;
	}
	if (acc == 123.456f) out[0] = acc;
}

int main() {
	const size_t bytes = (size_t)5000000 * 512;
	float4 *x; float *out;
	cudaMalloc(&x, bytes); cudaMalloc(&out, 4);
	cudaMemset(x, 0, bytes);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	float ms;
	#define RUN(name, launch) for (int r = 0; r < 3; r++) { cudaEventRecord(e0); launch; cudaEventRecord(e1); cudaEventSynchronize(e1); } \
		cudaEventElapsedTime(&ms, e0, e1); printf("%-40s %.3f ms  %.0f GB/s\n", name, ms, bytes / ms / 1e6);
	RUN("stride unr4 148x8x256", (read_stride<4><<<148 * 8, 256>>>(x, bytes / 16, out)))
	RUN("stride unr8 148x8x256", (read_stride<8><<<148 * 8, 256>>>(x, bytes / 16, out)))
	RUN("stride unr8 148x4x512", (read_stride<8><<<148 * 4, 512>>>(x, bytes / 16, out)))
	RUN("stride unr16 148x2x512", (read_stride<16><<<148 * 2, 512>>>(x, bytes / 16, out)))
	RUN("tiles 148x256", (read_tiles<<<148, 256>>>(x, (int)(bytes / 65536), out)))
	RUN("tiles 296x256", (read_tiles<<<296, 256>>>(x, (int)(bytes / 65536), out)))
	RUN("tiles 592x256", (read_tiles<<<592, 256>>>(x, (int)(bytes / 65536), out)))
	printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
	return 0;
}
