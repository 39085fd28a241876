// Multi-GPU helper of the fused gradient exchange. sm_100a, NVLink 5 / NVSwitch (NVLS multicast).
//
// The reference is single-GPU (utils/general_utils.py:133 pins cuda:0); SURVEY.md 8e defines the
// exchange of the replicated trainer: the SUM of the parameter gradients over the ranks. With the
// gradient buffer in symmetric memory and its multicast alias known, the gradient accumulation
// kernels issue their reductions as multimem.red (preprocess.cu, red_add1 / red_add4), so the
// all-reduce happens inside the producing kernel and in the switch. This file holds the same
// operation for small vectors that are not produced by those kernels (the gradient of the event
// threshold c).
#include "common.cuh"
#include "../../include/event3dgs.h"

namespace {

__global__ void multimem_add_kernel(float* __restrict__ mc_dst, const float* __restrict__ src, size_t n, float scale)
{
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
		const float v = src[i] * scale;
		if (v != 0.f) asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" :: "l"(mc_dst + i), "f"(v) : "memory");
	}
}

// In-switch all-reduce of a symmetric buffer through its multicast alias. Every rank owns 1/world of the
// 16-byte vectors: multimem.ld_reduce (LDGMC.E.ADD.F32x4) fetches the SUM of that vector over all ranks'
// copies, reduced inside the NVSwitch, and multimem.st writes scale * sum back into every copy. Per GPU
// this moves n/world in and n/world out plus (world-1) * n/world of incoming broadcasts, against
// 2 * n * (world-1)/world for a ring. The caller places a cross-rank barrier before (all local
// accumulation finished) and after (all copies written).
__device__ __forceinline__ float4 mc_ld_reduce(const float4* p)
{
	float4 v;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
		: "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void mc_st(float4* p, float4 v, float scale)
{
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(v.x * scale), "f"(v.y * scale),
		"f"(v.z * scale), "f"(v.w * scale) : "memory");
}

template <int U>
__global__ void __launch_bounds__(512) nvls_allreduce_kernel(float4* __restrict__ mc, size_t v_begin, size_t v_end, float scale)
{
	// U independent in-switch reductions in flight per thread before the first store
	const size_t stride = (size_t)gridDim.x * blockDim.x;
	size_t i = v_begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	for (; i + (U - 1) * stride < v_end; i += U * stride) {
		float4 v[U];
#pragma unroll
		for (int k = 0; k < U; k++) v[k] = mc_ld_reduce(mc + i + k * stride);
#pragma unroll
		for (int k = 0; k < U; k++) mc_st(mc + i + k * stride, v[k], scale);
	}
	for (; i < v_end; i += stride) mc_st(mc + i, mc_ld_reduce(mc + i), scale);
}

}  // namespace

extern "C" int e3_nvls_allreduce(float* multicast_ptr, size_t n, int rank, int world, float scale, void* stream)
{
	if (world < 1 || rank < 0 || rank >= world || (n & 3) || (reinterpret_cast<size_t>(multicast_ptr) & 15)) return E3_ERR_BAD_ARGUMENT;
	if (n == 0) return 0;
	if (!multicast_ptr) return E3_ERR_BAD_ARGUMENT;
	const size_t nv = n / 4;
	const size_t per = (nv + world - 1) / world;
	const size_t b = per * (size_t)rank < nv ? per * (size_t)rank : nv;
	const size_t e = b + per < nv ? b + per : nv;
	if (e > b) {
		// grid and unroll do not matter between 32..592 blocks and 4..8 vectors in flight per thread: the
		// kernel runs at what the fabric delivers (profiles/r1_nvls_allreduce_sweep_n2.txt)
		const int cfg_blocks = 148 * 4, cfg_unroll = 4;
		const size_t want = (e - b + 511) / 512;
		const unsigned blocks = (unsigned)(want < (size_t)cfg_blocks ? want : (size_t)cfg_blocks);
		float4* p = reinterpret_cast<float4*>(multicast_ptr);
		if (cfg_unroll >= 8) nvls_allreduce_kernel<8><<<blocks, 512, 0, (cudaStream_t)stream>>>(p, b, e, scale);
		else if (cfg_unroll >= 4) nvls_allreduce_kernel<4><<<blocks, 512, 0, (cudaStream_t)stream>>>(p, b, e, scale);
		else nvls_allreduce_kernel<1><<<blocks, 512, 0, (cudaStream_t)stream>>>(p, b, e, scale);
		g_e3_launches += 1;
	}
	return (int)cudaGetLastError();
}

extern "C" int e3_multimem_add(float* multicast_dst, const float* src, size_t n, float scale, void* stream)
{
	if (n == 0) return 0;
	if (!multicast_dst || !src) return E3_ERR_BAD_ARGUMENT;
	const unsigned blocks = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
	multimem_add_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(multicast_dst, src, n, scale);
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}
