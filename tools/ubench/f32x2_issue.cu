// Does a packed fp32x2 instruction (FFMA2 / FADD2 / FMUL2, sm_100) free issue slots? Three loops, each also
// carrying 64 integer (ALU pipe) instructions per iteration so that the loop is bound by issue slots:
//   A: 64 FFMA     B: 32 FFMA2 (the same 64 FMAs)     C: 32 FFMA (half the FMAs: the floor for B)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_issue f32x2_issue.cu && ./f32x2_issue
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, unsigned* outi, int iters, float s)
{
	float a[16], b[16];
	unsigned u[8];
#pragma unroll
	for (int i = 0; i < 16; i++) { a[i] = threadIdx.x * 0.001f + i; b[i] = i * 0.5f; }
#pragma unroll
	for (int i = 0; i < 8; i++) u[i] = threadIdx.x + i;
	for (int it = 0; it < iters; it++) {
#pragma unroll
		for (int r = 0; r < 4; r++) {
			if (MODE == 0) {
#pragma unroll
				for (int i = 0; i < 16; i++) a[i] = fmaf(a[i], s, b[i]);
			} else if (MODE == 1) {
#pragma unroll
				for (int i = 0; i < 16; i += 2)
					asm volatile("{.reg .b64 x, y, z; mov.b64 x, {%0,%1}; mov.b64 y, {%2,%2}; mov.b64 z, {%3,%4}; fma.rn.f32x2 x, x, y, z; mov.b64 {%0,%1}, x;}"
						: "+f"(a[i]), "+f"(a[i + 1]) : "f"(s), "f"(b[i]), "f"(b[i + 1]));
			} else {
#pragma unroll
				for (int i = 0; i < 8; i++) a[i] = fmaf(a[i], s, b[i]);
			}
#pragma unroll
			for (int i = 0; i < 8; i++) { u[i] = (u[i] ^ (u[(i + 1) & 7] >> 3)); u[i] = u[i] & 0x7fffffffu | 5u; }
		}
	}
	float acc = 0; unsigned ai = 0;
#pragma unroll
	for (int i = 0; i < 16; i++) acc += a[i];
#pragma unroll
	for (int i = 0; i < 8; i++) ai ^= u[i];
	out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
	outi[blockIdx.x * blockDim.x + threadIdx.x] = ai;
}

int main()
{
	float* o; unsigned* oi;
	cudaMalloc(&o, 148 * 8 * 256 * 4); cudaMalloc(&oi, 148 * 8 * 256 * 4);
	cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
	const int iters = 20000;
	for (int mode = 0; mode < 3; mode++) {
		float best = 1e9f;
		for (int rep = 0; rep < 4; rep++) {
			cudaEventRecord(e0);
			if (mode == 0) k<0><<<148 * 8, 256>>>(o, oi, iters, 0.999f);
			else if (mode == 1) k<1><<<148 * 8, 256>>>(o, oi, iters, 0.999f);
			else k<2><<<148 * 8, 256>>>(o, oi, iters, 0.999f);
			cudaEventRecord(e1); cudaEventSynchronize(e1);
			float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
		}
		printf("{\"mode\": \"%s\", \"ms\": %.3f}\n", mode == 0 ? "64 FFMA + 64 ALU" : mode == 1 ? "32 FFMA2 + 64 ALU" : "32 FFMA + 64 ALU", best);
	}
	printf("%s\n", cudaGetErrorString(cudaGetLastError()));
	return 0;
}
