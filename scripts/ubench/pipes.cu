// Micro-benchmark of the fp32 add / max pipes on sm_100a: which instruction mix bounds the WSB DP.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on a B200.
#include <cstdio>
#include <cuda_runtime.h>

#define N_CHAIN 32
#define ITERS 2048

__constant__ float c_w[64];
__constant__ float2 c_w2[32];

template<int MODE>
__global__ void __launch_bounds__(128) bench(float *out, float seed, long long *cycles) {
	float x[N_CHAIN];
	#pragma unroll
	for (int i = 0; i < N_CHAIN; i++) {
		x[i] = seed * (float)(i + threadIdx.x);
	}
	float y = seed * 3.0f, z = seed * 5.0f;
	long long t0 = clock64();
	#pragma unroll 1
	for (int it = 0; it < ITERS; it++) {
		if (MODE == 0) {          // FADD reg,reg
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) asm("add.rn.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y));
		} else if (MODE == 1) {   // FADD reg,const
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) x[i] = __fadd_rn(x[i], c_w[i]);
		} else if (MODE == 2) {   // FMNMX
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) asm("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(y));
		} else if (MODE == 3) {   // FMNMX3
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(y), "f"(z));
		} else if (MODE == 4) {   // add.f32x2
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i += 2) {
				asm("{ .reg .b64 a, b; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %3}; add.rn.f32x2 a, a, b; mov.b64 {%0, %1}, a; }"
					: "+f"(x[i]), "+f"(x[i + 1]) : "f"(y), "f"(z));
			}
		} else if (MODE == 5) {   // FADD + FMNMX per candidate (2 candidates per chain per iter)
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) {
				float a = __fadd_rn(y, c_w[i]), b = __fadd_rn(z, c_w[i + 1]);
				asm("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(a));
				asm("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(b));
				y += 1.0f;
			}
		} else if (MODE == 6) {   // 2 FADD + FMNMX3 per 2 candidates
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) {
				float a, b;
				asm("add.rn.f32 %0, %1, %2;" : "=f"(a) : "f"(x[(i + 1) % N_CHAIN]), "f"(y));
				asm("add.rn.f32 %0, %1, %2;" : "=f"(b) : "f"(x[(i + 2) % N_CHAIN]), "f"(z));
				asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
			}
		} else if (MODE == 7) {   // FADD2 + FMNMX3 per 2 candidates
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) {
				float a, b;
				const int p = ((i / 2 + 1) % (N_CHAIN / 2)) * 2;
				if (i & 1) {
					asm("{ .reg .b64 p, q; mov.b64 p, {%2, %3}; mov.b64 q, {%4, %5}; add.rn.f32x2 p, p, q; mov.b64 {%0, %1}, p; }"
						: "=f"(a), "=f"(b) : "f"(x[p]), "f"(x[p + 1]), "f"(y), "f"(z));
				} else {
					asm("{ .reg .b64 p, q; mov.b64 p, {%2, %3}; mov.b64 q, {%4, %5}; add.rn.f32x2 p, p, q; mov.b64 {%0, %1}, p; }"
						: "=f"(a), "=f"(b) : "f"(x[p]), "f"(x[p + 1]), "f"(z), "f"(y));
				}
				asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
			}
		} else if (MODE == 8) {   // VIMNMX3
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) {
				int xi = __float_as_int(x[i]);
				xi = max(max(xi, __float_as_int(y) + i), __float_as_int(z) + it);
				x[i] = __int_as_float(xi);
			}
		} else if (MODE == 10) {  // FADD2 r, UR pair (gap-cost pair from the parameter bank, as in the WSB kernel)
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i += 2) {
				const float2 w = c_w2[(i / 2 + it) & 31];
				asm("{ .reg .b64 a, b; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %3}; add.rn.f32x2 a, a, b; mov.b64 {%0, %1}, a; }"
					: "+f"(x[i]), "+f"(x[i + 1]) : "f"(w.x), "f"(w.y));
			}
		} else if (MODE == 11) {  // the WSB inner loop after the rewrite: 2 FADD2 (r, UR) + 2 FMNMX3 per 4 candidates
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i += 2) {
				const int p = (i + 2) % N_CHAIN, q = (i + 4) % N_CHAIN;
				const float2 w0 = c_w2[(i / 2 + it) & 31], w1 = c_w2[(i / 2 + it + 7) & 31];
				float a0, a1, b0, b1;
				asm("{ .reg .b64 a, b; mov.b64 a, {%2, %3}; mov.b64 b, {%4, %5}; add.rn.f32x2 a, a, b; mov.b64 {%0, %1}, a; }"
					: "=f"(a0), "=f"(a1) : "f"(x[p]), "f"(x[p + 1]), "f"(w0.x), "f"(w0.y));
				asm("{ .reg .b64 a, b; mov.b64 a, {%2, %3}; mov.b64 b, {%4, %5}; add.rn.f32x2 a, a, b; mov.b64 {%0, %1}, a; }"
					: "=f"(b0), "=f"(b1) : "f"(x[q]), "f"(x[q + 1]), "f"(w1.x), "f"(w1.y));
				asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a0), "f"(b0));
				asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i + 1]) : "f"(a1), "f"(b1));
			}
		} else if (MODE == 12) {  // FFMA2 for reference (fma.rn.f32x2)
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i += 2) {
				asm("{ .reg .b64 a, b; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %3}; fma.rn.f32x2 a, a, b, b; mov.b64 {%0, %1}, a; }"
					: "+f"(x[i]), "+f"(x[i + 1]) : "f"(y), "f"(z));
			}
		} else if (MODE == 9) {   // 2 FADD(reg,const) + FMNMX3
			#pragma unroll
			for (int i = 0; i < N_CHAIN; i++) {
				float a = __fadd_rn(x[(i + 1) % N_CHAIN], c_w[i]);
				float b = __fadd_rn(x[(i + 2) % N_CHAIN], c_w[i + 1]);
				asm("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
			}
		}
	}
	long long t1 = clock64();
	float s = 0;
	#pragma unroll
	for (int i = 0; i < N_CHAIN; i++) s += x[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = s + y + z;
	if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template<int MODE>
void run(const char *name, int instr_per_iter, int cand_per_iter, int warps_per_sm) {
	int sms = 148;
	float *out; long long *cyc;
	int blocks = sms * (warps_per_sm / 4);
	cudaMalloc(&out, sizeof(float) * blocks * 128);
	cudaMalloc(&cyc, sizeof(long long) * blocks);
	bench<MODE><<<blocks, 128>>>(out, 1e-3f, cyc);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	cudaEventRecord(e0);
	bench<MODE><<<blocks, 128>>>(out, 1e-3f, cyc);
	cudaEventRecord(e1);
	cudaDeviceSynchronize();
	float ms; cudaEventElapsedTime(&ms, e0, e1);
	long long h[4096]; cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
	double avg = 0; for (int i = 0; i < blocks; i++) avg += h[i]; avg /= blocks;
	double warps_per_smsp = warps_per_sm / 4.0;
	double cyc_per_instr = avg / ((double)ITERS * instr_per_iter * warps_per_smsp);
	printf("%-34s warps/SM %2d  cyc/warp-instr/SMSP %.3f  cyc/candidate-warp/SMSP %.3f  (%.3f ms)\n", name, warps_per_sm,
		cyc_per_instr, cand_per_iter ? avg / ((double)ITERS * cand_per_iter * warps_per_smsp) : 0.0, ms);
	cudaFree(out); cudaFree(cyc);
}

int main() {
	float w[64]; for (int i = 0; i < 64; i++) w[i] = 0.01f * i;
	cudaMemcpyToSymbol(c_w, w, sizeof(w));
	float2 w2[32]; for (int i = 0; i < 32; i++) w2[i] = make_float2(-0.01f * i, -0.01f * i);
	cudaMemcpyToSymbol(c_w2, w2, sizeof(w2));
	for (int wps : {4, 8, 12, 16}) {
		run<0>("FADD r,r", N_CHAIN, 0, wps);
		run<1>("FADD r,c", N_CHAIN, 0, wps);
		run<2>("FMNMX", N_CHAIN, 0, wps);
		run<3>("FMNMX3", N_CHAIN, 0, wps);
		run<4>("FADD2 (add.f32x2)", N_CHAIN / 2, 0, wps);
		run<5>("2x(FADD r,c + FMNMX) [+1 FADD]", N_CHAIN * 5, N_CHAIN * 2, wps);
		run<6>("2 FADD r,r + FMNMX3", N_CHAIN * 3, N_CHAIN * 2, wps);
		run<7>("FADD2 + FMNMX3", N_CHAIN * 2, N_CHAIN * 2, wps);
		run<8>("VIMNMX3 (+2 IADD)", N_CHAIN * 3, 0, wps);
		run<9>("2 FADD r,c + FMNMX3", N_CHAIN * 3, N_CHAIN * 2, wps);
		run<10>("FADD2 r,UR", N_CHAIN / 2, 0, wps);
		run<11>("2 FADD2 r,UR + 2 FMNMX3 (4 cand)", N_CHAIN * 2, N_CHAIN * 2, wps);
		run<12>("FFMA2", N_CHAIN / 2, 0, wps);
	}
	return 0;
}
