// TMEM as a per-thread scratchpad: correctness of tcgen05.alloc/st/ld 32x32b and its load throughput/latency.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem tmem.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld4(uint32_t taddr, float (&r)[4]) {
	uint32_t a, b, c, d;
	asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
		: "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(taddr));
	r[0] = __uint_as_float(a); r[1] = __uint_as_float(b); r[2] = __uint_as_float(c); r[3] = __uint_as_float(d);
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, const float (&r)[4]) {
	asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
		:: "r"(taddr), "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])), "r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__global__ void __launch_bounds__(256, 1) tmem_test(int *errors, long long *cycles, float *sink, int iters) {
	__shared__ uint32_t tmem_base_sh;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	if (warp == 0) {
		uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base_sh);
		uint32_t ncols = 512;
		asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(dst), "r"(ncols) : "memory");
		asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
	}
	asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
	__syncthreads();
	asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
	const uint32_t base = tmem_base_sh;
	// warp w owns lanes 32*(w%4) .. +31 and columns 256*(w/4) .. +255
	const uint32_t mine = base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)(256 * (warp >> 2));

	int bad = 0;
	for (int c = 0; c < 256; c += 4) {
		float v[4];
		for (int j = 0; j < 4; j++) v[j] = (float)(threadIdx.x * 1000 + c + j);
		tmem_st4(mine + c, v);
	}
	tmem_wait_st();
	for (int c = 0; c < 256; c += 4) {
		float v[4];
		tmem_ld4(mine + c, v);
		tmem_wait_ld();
		for (int j = 0; j < 4; j++) bad += v[j] != (float)(threadIdx.x * 1000 + c + j);
	}
	atomicAdd(errors, bad);
	__syncthreads();

	// throughput: every warp streams its 252 columns (21 rows of 12) repeatedly
	float acc = 0.0f;
	long long t0 = clock64();
	for (int it = 0; it < iters; it++) {
		#pragma unroll 1
		for (int row = 0; row < 21; row++) {
			float a[4], b[4], c[4];
			tmem_ld4(mine + row * 12, a);
			tmem_ld4(mine + row * 12 + 4, b);
			tmem_ld4(mine + row * 12 + 8, c);
			tmem_wait_ld();
			acc += a[0] + b[1] + c[2];
		}
	}
	long long t1 = clock64();
	// latency: dependent single loads
	long long t2 = clock64();
	uint32_t col = 0;
	for (int it = 0; it < 256; it++) {
		float a[4];
		tmem_ld4(mine + col, a);
		tmem_wait_ld();
		col = ((uint32_t)a[3]) & 0xfc & 0x7f;   // data-dependent next column
	}
	long long t3 = clock64();
	sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + (float)col;
	if (lane == 0) {
		cycles[(blockIdx.x * 8 + warp) * 2] = t1 - t0;
		cycles[(blockIdx.x * 8 + warp) * 2 + 1] = t3 - t2;
	}
	__syncthreads();
	if (warp == 0) {
		uint32_t ncols = 512;
		asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(base), "r"(ncols) : "memory");
	}
}

int main() {
	int *err; long long *cyc; float *sink;
	const int blocks = 148, iters = 200;
	cudaMalloc(&err, 4); cudaMemset(err, 0, 4);
	cudaMalloc(&cyc, sizeof(long long) * blocks * 16);
	cudaMalloc(&sink, sizeof(float) * blocks * 256);
	tmem_test<<<blocks, 256>>>(err, cyc, sink, iters);
	cudaError_t e = cudaDeviceSynchronize();
	printf("status: %s\n", cudaGetErrorString(e));
	int h_err = -1; cudaMemcpy(&h_err, err, 4, cudaMemcpyDeviceToHost);
	long long h[148 * 16]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
	double a = 0, b = 0; for (int i = 0; i < blocks * 8; i++) { a += h[2 * i]; b += h[2 * i + 1]; }
	a /= blocks * 8; b /= blocks * 8;
	printf("errors %d\n", h_err);
	double bytes_per_warp = (double)iters * 21 * 12 * 32 * 4;
	printf("stream: %.0f cycles per warp; SM-wide %.1f B/cycle (8 warps)\n", a, 8 * bytes_per_warp / a);
	printf("dependent ld+wait latency: %.1f cycles\n", b / 256);
	// second launch to check that alloc/dealloc left the SM usable
	cudaMemset(err, 0, 4);
	tmem_test<<<blocks * 3, 256>>>(err, cyc, sink, 10);
	e = cudaDeviceSynchronize();
	cudaMemcpy(&h_err, err, 4, cudaMemcpyDeviceToHost);
	printf("second launch (3 waves): %s errors %d\n", cudaGetErrorString(e), h_err);
	return 0;
}
