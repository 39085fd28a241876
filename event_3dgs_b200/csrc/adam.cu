// Fused Adam step over all parameter groups in ONE launch. sm_100a.
//
// Reference: scene/gaussian_model.py:149-175 builds torch.optim.Adam(lr=0, eps=1e-15) with six
// parameter groups (xyz, f_dc, f_rest, opacity, scaling, rotation; 59 floats per Gaussian) and
// train.py:330-332 steps it every iteration; train.py:71-73 keeps a second Adam for the scalar
// event threshold c. torch runs several element-wise kernels per group. The arithmetic here is
// torch's (no amsgrad, no weight decay):
//   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// Algorithmic bytes: read p, g, m, v and write p, m, v = 28 B per parameter (1652 B per
// Gaussian): a pure HBM stream, 16-byte vectorised.
#include "common.cuh"
#include "../../include/event3dgs.h"

namespace {

struct AdamTable {
	float* p[E3_ADAM_MAX_GROUPS];
	const float* g[E3_ADAM_MAX_GROUPS];
	float* m[E3_ADAM_MAX_GROUPS];
	float* v[E3_ADAM_MAX_GROUPS];
	unsigned long long n[E3_ADAM_MAX_GROUPS];
	unsigned first_block[E3_ADAM_MAX_GROUPS + 1];
	float step_size[E3_ADAM_MAX_GROUPS];     // lr / (1 - b1^t)
	float eps[E3_ADAM_MAX_GROUPS];
	int num;
};

constexpr int kAdamThreads = 256;
constexpr int kAdamPerBlock = kAdamThreads * 8;     // two float4 per thread

// omb1 / omb2 = 1 - beta, rounded from double like torch does (1 - 0.999 in fp32 arithmetic would be
// off by 1e-5 relative)
__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, float b1, float b2, float omb1, float omb2,
	float step_size, float inv_bc2_sqrt, float eps)
{
	m = b1 * m + omb1 * g;
	v = b2 * v + omb2 * g * g;
	const float denom = sqrtf(v) * inv_bc2_sqrt + eps;
	p = p - step_size * (m / denom);
}

__global__ void __launch_bounds__(kAdamThreads) adam_kernel(const AdamTable tab, float b1, float b2, float omb1, float omb2,
	float inv_bc2_sqrt)
{
	int gi = 0;
#pragma unroll
	for (int i = 1; i < E3_ADAM_MAX_GROUPS; i++)
		if (i < tab.num && blockIdx.x >= tab.first_block[i]) gi = i;
	float* __restrict__ P = tab.p[gi];
	const float* __restrict__ G = tab.g[gi];
	float* __restrict__ M = tab.m[gi];
	float* __restrict__ V = tab.v[gi];
	const unsigned long long n = tab.n[gi];
	const float ss = tab.step_size[gi], eps = tab.eps[gi];
	const unsigned long long base = (unsigned long long)(blockIdx.x - tab.first_block[gi]) * kAdamPerBlock;
	const bool aligned = ((reinterpret_cast<size_t>(P) | reinterpret_cast<size_t>(G) | reinterpret_cast<size_t>(M) |
		reinterpret_cast<size_t>(V)) & 15) == 0;
#pragma unroll
	for (int h = 0; h < 2; h++) {
		const unsigned long long i = base + ((unsigned long long)h * kAdamThreads + threadIdx.x) * 4;
		if (aligned && i + 3 < n) {
			float4 p = *reinterpret_cast<float4*>(P + i), m = *reinterpret_cast<float4*>(M + i), v = *reinterpret_cast<float4*>(V + i);
			const float4 g = __ldg(reinterpret_cast<const float4*>(G + i));
			adam1(p.x, g.x, m.x, v.x, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
			adam1(p.y, g.y, m.y, v.y, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
			adam1(p.z, g.z, m.z, v.z, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
			adam1(p.w, g.w, m.w, v.w, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
			*reinterpret_cast<float4*>(P + i) = p; *reinterpret_cast<float4*>(M + i) = m; *reinterpret_cast<float4*>(V + i) = v;
		} else {
			for (int k = 0; k < 4; k++)
				if (i + k < n) {
					float p = P[i + k], m = M[i + k], v = V[i + k];
					adam1(p, G[i + k], m, v, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
					P[i + k] = p; M[i + k] = m; V[i + k] = v;
				}
		}
	}
}

// ---------------------------------------------------------------------------------------------
// Sharded Adam fused with the gradient exchange (upgrade path of the replicated
// multi-GPU trainer). Gradients and parameters live in symmetric memory with NVLS multicast
// aliases. Rank r owns vectors [r * per, (r+1) * per) of every group and, in ONE kernel,
//   multimem.ld_reduce  the SUM over all ranks of its gradient vectors (reduced in the NVSwitch),
//   Adam                on those rows only (its 1/world of the moments),
//   multimem.st         the new parameters into EVERY rank's copy.
// Against all-reduce + replicated Adam this removes the all-reduce's store phase and (world-1)/world
// of the optimizer's 28 bytes per parameter. The moments of rows a rank does not own go stale on it;
// the host gathers them when it needs full moments (densification, checkpoints).
struct ShardTable {
	float* p[E3_ADAM_MAX_GROUPS];
	const float* g[E3_ADAM_MAX_GROUPS];
	float* m[E3_ADAM_MAX_GROUPS];
	float* v[E3_ADAM_MAX_GROUPS];
	unsigned long long n[E3_ADAM_MAX_GROUPS];          // floats in the group (moments hold exactly n)
	unsigned long long v_begin[E3_ADAM_MAX_GROUPS];    // first owned float4 vector
	unsigned long long v_end[E3_ADAM_MAX_GROUPS];
	unsigned first_block[E3_ADAM_MAX_GROUPS + 1];
	float step_size[E3_ADAM_MAX_GROUPS];
	float eps[E3_ADAM_MAX_GROUPS];
	int num;
};

__global__ void __launch_bounds__(kAdamThreads) sharded_adam_kernel(const ShardTable tab, float b1, float b2, float omb1, float omb2,
	float inv_bc2_sqrt, long long grad_mc_delta, long long param_mc_delta, float grad_scale)
{
	int gi = 0;
#pragma unroll
	for (int i = 1; i < E3_ADAM_MAX_GROUPS; i++)
		if (i < tab.num && blockIdx.x >= tab.first_block[i]) gi = i;
	float* __restrict__ P = tab.p[gi];
	float* __restrict__ M = tab.m[gi];
	float* __restrict__ V = tab.v[gi];
	const float4* g_mc = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(tab.g[gi]) + grad_mc_delta);
	float4* p_mc = reinterpret_cast<float4*>(reinterpret_cast<char*>(P) + param_mc_delta);
	const unsigned long long n = tab.n[gi];
	const float ss = tab.step_size[gi], eps = tab.eps[gi];
	const unsigned long long i = tab.v_begin[gi] + (unsigned long long)(blockIdx.x - tab.first_block[gi]) * kAdamThreads + threadIdx.x;
	if (i >= tab.v_end[gi]) return;
	float4 g;
	asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
		: "=f"(g.x), "=f"(g.y), "=f"(g.z), "=f"(g.w) : "l"(g_mc + i) : "memory");
	g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
	float4 p = reinterpret_cast<const float4*>(P)[i];          // parameter segments are padded to whole vectors
	if (4 * i + 3 < n) {
		float4 m = reinterpret_cast<float4*>(M)[i], v = reinterpret_cast<float4*>(V)[i];
		adam1(p.x, g.x, m.x, v.x, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
		adam1(p.y, g.y, m.y, v.y, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
		adam1(p.z, g.z, m.z, v.z, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
		adam1(p.w, g.w, m.w, v.w, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
		reinterpret_cast<float4*>(M)[i] = m; reinterpret_cast<float4*>(V)[i] = v;
	} else {
		// last, partial vector of the group: the moment tensors end at n
		float pv[4] = { p.x, p.y, p.z, p.w };
		const float gv[4] = { g.x, g.y, g.z, g.w };
		for (int k = 0; k < 4; k++)
			if (4 * i + k < n) {
				float m = M[4 * i + k], v = V[4 * i + k];
				adam1(pv[k], gv[k], m, v, b1, b2, omb1, omb2, ss, inv_bc2_sqrt, eps);
				M[4 * i + k] = m; V[4 * i + k] = v;
			}
		p = make_float4(pv[0], pv[1], pv[2], pv[3]);
	}
	asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p_mc + i), "f"(p.x), "f"(p.y), "f"(p.z), "f"(p.w) : "memory");
}

}  // namespace

extern "C" int e3_nvls_sharded_adam(const e3_adam_group* groups_host, int num_groups, int step, double beta1, double beta2,
	long long grad_multicast_delta, long long param_multicast_delta, int rank, int world, float grad_scale, void* stream)
{
	if (num_groups < 0 || num_groups > E3_ADAM_MAX_GROUPS || step < 1 || (num_groups > 0 && !groups_host) || world < 1 || rank < 0 ||
		rank >= world || grad_multicast_delta == 0 || param_multicast_delta == 0)
		return E3_ERR_BAD_ARGUMENT;
	ShardTable tab;
	tab.num = 0;
	unsigned blocks = 0;
	const double bc1 = 1.0 - pow(beta1, (double)step);
	const double bc2 = 1.0 - pow(beta2, (double)step);
	for (int i = 0; i < num_groups; i++) {
		const e3_adam_group& g = groups_host[i];
		if (g.n == 0) continue;
		if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) return E3_ERR_BAD_ARGUMENT;
		if ((reinterpret_cast<size_t>(g.param) | reinterpret_cast<size_t>(g.grad) | reinterpret_cast<size_t>(g.exp_avg) |
			reinterpret_cast<size_t>(g.exp_avg_sq)) & 15)
			return E3_ERR_BAD_ARGUMENT;
		const unsigned long long nv = (g.n + 3) / 4, per = (nv + world - 1) / world;
		const unsigned long long b = per * rank < nv ? per * rank : nv, e = b + per < nv ? b + per : nv;
		if (e <= b) continue;
		const int k = tab.num++;
		tab.p[k] = g.param; tab.g[k] = g.grad; tab.m[k] = g.exp_avg; tab.v[k] = g.exp_avg_sq;
		tab.n[k] = g.n; tab.v_begin[k] = b; tab.v_end[k] = e; tab.first_block[k] = blocks;
		tab.step_size[k] = (float)((double)g.lr / bc1);
		tab.eps[k] = g.eps;
		blocks += (unsigned)((e - b + kAdamThreads - 1) / kAdamThreads);
	}
	if (tab.num == 0) return 0;
	tab.first_block[tab.num] = blocks;
	sharded_adam_kernel<<<blocks, kAdamThreads, 0, (cudaStream_t)stream>>>(tab, (float)beta1, (float)beta2, (float)(1.0 - beta1),
		(float)(1.0 - beta2), (float)(1.0 / sqrt(bc2)), grad_multicast_delta, param_multicast_delta, grad_scale);
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}

extern "C" int e3_adam_step(const e3_adam_group* groups_host, int num_groups, int step, double beta1, double beta2, void* stream)
{
	if (num_groups < 0 || num_groups > E3_ADAM_MAX_GROUPS || step < 1 || (num_groups > 0 && !groups_host)) return E3_ERR_BAD_ARGUMENT;
	AdamTable tab;
	tab.num = 0;
	unsigned blocks = 0;
	const double bc1 = 1.0 - pow(beta1, (double)step);
	const double bc2 = 1.0 - pow(beta2, (double)step);
	for (int i = 0; i < num_groups; i++) {
		const e3_adam_group& g = groups_host[i];
		if (g.n == 0) continue;
		if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) return E3_ERR_BAD_ARGUMENT;
		const int k = tab.num++;
		tab.p[k] = g.param; tab.g[k] = g.grad; tab.m[k] = g.exp_avg; tab.v[k] = g.exp_avg_sq;
		tab.n[k] = g.n; tab.first_block[k] = blocks;
		tab.step_size[k] = (float)((double)g.lr / bc1);
		tab.eps[k] = g.eps;
		blocks += (unsigned)((g.n + kAdamPerBlock - 1) / kAdamPerBlock);
	}
	if (tab.num == 0) return 0;
	tab.first_block[tab.num] = blocks;
	adam_kernel<<<blocks, kAdamThreads, 0, (cudaStream_t)stream>>>(tab, (float)beta1, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2),
		(float)(1.0 / sqrt(bc2)));
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}
