// Adaptive density control (clone / split / prune) with optimizer-state surgery, and the
// per-iteration densification statistics. sm_100a.
//
// Reference: scene/gaussian_model.py:258-407 (replace_tensor_to_optimizer, _prune_optimizer,
// prune_points, cat_tensors_to_optimizer, densification_postfix, densify_and_split,
// densify_and_clone, densify_and_prune, add_densification_stats), reset_opacity :210-213,
// schedule train.py:317-327. The reference runs this as ~150 ATen launches: boolean-mask indexing
// and torch.cat of every parameter and of both Adam moments, three times over (after clone,
// after split, after prune), each allocating. The net effect is a stable compaction into four
// segments:
//   [ kept originals | kept clones | kept split children, 1st draw | kept split children, 2nd draw ]
// so it is done here as: one flag kernel, four prefix sums, one row-map kernel, then ONE gather
// launch over all arrays (parameters and moments), reading 708 B and writing <= 708 B per kept row.
//
//   original i is cloned  iff |g_i| >= max_grad and max exp(scaling_i) <= percent_dense * extent
//   original i is split   iff  g_i  >= max_grad and max exp(scaling_i) >  percent_dense * extent
//       (g = xyz_gradient_accum / denom with NaN -> 0; a split original is removed, its two
//        children get scaling log(exp(s) / 1.6) and xyz + R(q) (exp(s) * z))
//   afterwards every row (old and new) is pruned iff sigmoid(opacity) < min_opacity, or (when a
//   screen-size limit is given) max exp(scaling) > 0.1 * extent. The reference's third criterion,
//   max_radii2D > max_screen_size, can never fire: densification_postfix has just zeroed
//   max_radii2D (gaussian_model.py:347), so it is not evaluated.
// All statistics of the surviving rows are zero afterwards (gaussian_model.py:345-347).
#include "common.cuh"
#include "../../include/event3dgs.h"

namespace {

constexpr int kThreads = 256;

// scratch layout (uint32 words)
struct DensifyScratch {
	uint32_t *flag_keep, *flag_clone, *flag_child, *flag_sel;   // [P] 0/1
	uint32_t *off_keep, *off_clone, *off_child, *rank_sel;      // [P] exclusive sums of the flags
	uint32_t* scan_tmp;                                         // [scan_blocks(P) + 2]
	uint32_t* row_map;                                          // [2P] source row | kind << 30
	static DensifyScratch carve(char* p, size_t P, char** end = nullptr)
	{
		DensifyScratch s;
		e3_carve(p, s.flag_keep, P); e3_carve(p, s.flag_clone, P); e3_carve(p, s.flag_child, P); e3_carve(p, s.flag_sel, P);
		e3_carve(p, s.off_keep, P); e3_carve(p, s.off_clone, P); e3_carve(p, s.off_child, P); e3_carve(p, s.rank_sel, P);
		e3_carve(p, s.scan_tmp, e3_scan_blocks(P) + 2);
		e3_carve(p, s.row_map, 2 * P);
		if (end) *end = p;
		return s;
	}
};

__device__ __forceinline__ float sigmoid_aten(float x) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x))); }

__global__ void __launch_bounds__(kThreads) densify_flags_kernel(int P, const float* __restrict__ accum,
	const float* __restrict__ denom, const float* __restrict__ scaling, const float* __restrict__ opacity, float max_grad,
	float dense_extent, float min_opacity, float big_ws, float inv_split, DensifyScratch s)
{
	const int i = blockIdx.x * kThreads + threadIdx.x;
	if (i >= P) return;
	float g = __fdiv_rn(accum[i], denom[i]);
	if (g != g) g = 0.0f;
	const float s0 = expf(scaling[3 * i]), s1 = expf(scaling[3 * i + 1]), s2 = expf(scaling[3 * i + 2]);
	const float smax = fmaxf(fmaxf(s0, s1), s2);
	const bool clone = fabsf(g) >= max_grad && smax <= dense_extent;
	const bool split = g >= max_grad && smax > dense_extent;
	const bool low_opacity = sigmoid_aten(opacity[i]) < min_opacity;
	bool prune_self = low_opacity, prune_child = low_opacity;
	if (big_ws >= 0.0f) {
		prune_self = prune_self || smax > big_ws;
		// children: get_scaling of the new raw scaling log(exp(s) * (1 / 1.6))
		const float c0 = expf(logf(__fmul_rn(s0, inv_split))), c1 = expf(logf(__fmul_rn(s1, inv_split))),
			c2 = expf(logf(__fmul_rn(s2, inv_split)));
		prune_child = prune_child || fmaxf(fmaxf(c0, c1), c2) > big_ws;
	}
	s.flag_keep[i] = (!split && !prune_self) ? 1u : 0u;
	s.flag_clone[i] = (clone && !prune_self) ? 1u : 0u;
	s.flag_child[i] = (split && !prune_child) ? 1u : 0u;
	s.flag_sel[i] = split ? 1u : 0u;
}

// counts: 0 kept originals, 1 kept clones, 2 kept children per draw, 3 split selections, 4 new row count
__global__ void __launch_bounds__(kThreads) densify_map_kernel(int P, DensifyScratch s, int* __restrict__ counts)
{
	const int i = blockIdx.x * kThreads + threadIdx.x;
	const uint32_t n_keep = (uint32_t)counts[0], n_clone = (uint32_t)counts[1], n_child = (uint32_t)counts[2];
	if (i == 0) counts[4] = (int)(n_keep + n_clone + 2 * n_child);
	if (i >= P) return;
	if (s.flag_keep[i]) s.row_map[s.off_keep[i]] = (uint32_t)i;
	if (s.flag_clone[i]) s.row_map[n_keep + s.off_clone[i]] = (uint32_t)i | (1u << 30);
	if (s.flag_child[i]) {
		s.row_map[n_keep + n_clone + s.off_child[i]] = (uint32_t)i | (2u << 30);
		s.row_map[n_keep + n_clone + n_child + s.off_child[i]] = (uint32_t)i | (3u << 30);
	}
}


struct GatherTable {
	const float* src[E3_DENSIFY_MAX_ARRAYS];
	float* dst[E3_DENSIFY_MAX_ARRAYS];
	const float* m_src[E3_DENSIFY_MAX_ARRAYS];
	float* m_dst[E3_DENSIFY_MAX_ARRAYS];
	const float* v_src[E3_DENSIFY_MAX_ARRAYS];
	float* v_dst[E3_DENSIFY_MAX_ARRAYS];
	int width[E3_DENSIFY_MAX_ARRAYS];
	int kind[E3_DENSIFY_MAX_ARRAYS];
	unsigned first_block[E3_DENSIFY_MAX_ARRAYS + 1];
	int num;
};

constexpr int kGatherPerThread = 4;
constexpr int kGatherPerBlock = kThreads * kGatherPerThread;

// rotation matrix row k of build_rotation (utils/general_utils.py:78-99), every operation rounded
// on its own like the chain of ATen element-wise kernels it replaces
__device__ __forceinline__ void rotation_row(const float* __restrict__ q4, int k, float& a, float& b, float& c)
{
	const float q0 = q4[0], q1 = q4[1], q2 = q4[2], q3 = q4[3];
	const float n = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, q0), __fmul_rn(q1, q1)), __fmul_rn(q2, q2)), __fmul_rn(q3, q3)));
	const float r = __fdiv_rn(q0, n), x = __fdiv_rn(q1, n), y = __fdiv_rn(q2, n), z = __fdiv_rn(q3, n);
	auto mul = [](float u, float v) { return __fmul_rn(u, v); };
	if (k == 0) {
		a = __fsub_rn(1.0f, mul(2.0f, __fadd_rn(mul(y, y), mul(z, z))));
		b = mul(2.0f, __fsub_rn(mul(x, y), mul(r, z)));
		c = mul(2.0f, __fadd_rn(mul(x, z), mul(r, y)));
	} else if (k == 1) {
		a = mul(2.0f, __fadd_rn(mul(x, y), mul(r, z)));
		b = __fsub_rn(1.0f, mul(2.0f, __fadd_rn(mul(x, x), mul(z, z))));
		c = mul(2.0f, __fsub_rn(mul(y, z), mul(r, x)));
	} else {
		a = mul(2.0f, __fsub_rn(mul(x, z), mul(r, y)));
		b = mul(2.0f, __fadd_rn(mul(y, z), mul(r, x)));
		c = __fsub_rn(1.0f, mul(2.0f, __fadd_rn(mul(x, x), mul(y, y))));
	}
}

__global__ void __launch_bounds__(kThreads) densify_gather_kernel(const GatherTable tab, unsigned long long rows,
	const uint32_t* __restrict__ row_map, const uint32_t* __restrict__ rank_sel, const float* __restrict__ rotation,
	const float* __restrict__ scaling, const float* __restrict__ normals, unsigned n_sel, float inv_split)
{
	int a = 0;
#pragma unroll
	for (int i = 1; i < E3_DENSIFY_MAX_ARRAYS; i++)
		if (i < tab.num && blockIdx.x >= tab.first_block[i]) a = i;
	const int w = tab.width[a], kind = tab.kind[a];
	const float* __restrict__ src = tab.src[a];
	float* __restrict__ dst = tab.dst[a];
	const float* __restrict__ m_src = tab.m_src[a];
	const float* __restrict__ v_src = tab.v_src[a];
	float* __restrict__ m_dst = tab.m_dst[a];
	float* __restrict__ v_dst = tab.v_dst[a];
	const unsigned long long total = rows * (unsigned long long)w;
	const unsigned long long base = (unsigned long long)(blockIdx.x - tab.first_block[a]) * kGatherPerBlock;
#pragma unroll
	for (int h = 0; h < kGatherPerThread; h++) {
		const unsigned long long e = base + (unsigned long long)h * kThreads + threadIdx.x;
		if (e >= total) break;
		const unsigned long long r = e / (unsigned)w;
		const int k = (int)(e - r * (unsigned)w);
		const uint32_t m = row_map[r];
		const uint32_t i = m & 0x3fffffffu, what = m >> 30;
		float val = src[(size_t)i * w + k];
		if (what >= 2) {
			if (kind == 1) {
				// new_xyz = R(q) (exp(s) * z) + xyz, gaussian_model.py:358-362
				float r0, r1, r2;
				rotation_row(rotation + 4 * (size_t)i, k, r0, r1, r2);
				const float* z = normals + 3 * ((size_t)(what - 2) * n_sel + rank_sel[i]);
				const float* sc = scaling + 3 * (size_t)i;
				const float d0 = __fmul_rn(z[0], expf(sc[0])), d1 = __fmul_rn(z[1], expf(sc[1])), d2 = __fmul_rn(z[2], expf(sc[2]));
				val = __fadd_rn(__fmaf_rn(r2, d2, __fmaf_rn(r1, d1, __fmul_rn(r0, d0))), val);
			} else if (kind == 2) {
				val = logf(__fmul_rn(expf(val), inv_split));   // gaussian_model.py:363
			}
		}
		dst[e] = val;
		if (m_dst) m_dst[e] = (what == 0 && m_src) ? m_src[(size_t)i * w + k] : 0.0f;
		if (v_dst) v_dst[e] = (what == 0 && v_src) ? v_src[(size_t)i * w + k] : 0.0f;
	}
}

__global__ void __launch_bounds__(kThreads) densification_stats_kernel(int P, const int* __restrict__ radii,
	const float* __restrict__ d_mean2D, float* __restrict__ max_radii2D, float* __restrict__ accum, float* __restrict__ denom)
{
	const int i = blockIdx.x * kThreads + threadIdx.x;
	if (i >= P) return;
	const int r = radii[i];
	if (r <= 0) return;
	max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
	const float gx = d_mean2D[3 * (size_t)i], gy = d_mean2D[3 * (size_t)i + 1];
	accum[i] = __fadd_rn(accum[i], __fsqrt_rn(__fadd_rn(__fmul_rn(gx, gx), __fmul_rn(gy, gy))));
	denom[i] = __fadd_rn(denom[i], 1.0f);
}

// reset_opacity, gaussian_model.py:210-213: inverse_sigmoid(min(sigmoid(o), 0.01)), moments zeroed
__global__ void __launch_bounds__(kThreads) reset_opacity_kernel(int P, float* __restrict__ opacity, float* __restrict__ m,
	float* __restrict__ v, float cap)
{
	const int i = blockIdx.x * kThreads + threadIdx.x;
	if (i >= P) return;
	const float x = fminf(sigmoid_aten(opacity[i]), cap);
	opacity[i] = logf(__fdiv_rn(x, __fsub_rn(1.0f, x)));
	if (m) m[i] = 0.0f;
	if (v) v[i] = 0.0f;
}

}  // namespace

#define DCHECK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return (int)e_; } while (0)

extern "C" size_t e3_densify_scratch_bytes(int P)
{
	if (P <= 0) return 256;
	char* end = nullptr;
	DensifyScratch::carve(nullptr, (size_t)P, &end);
	return reinterpret_cast<size_t>(end) + 256;
}

extern "C" int e3_densify_plan(int P, const float* xyz_gradient_accum, const float* denom, const float* scaling,
	const float* opacity, float max_grad, float dense_extent, float min_opacity, float big_world_extent, int split_n,
	void* scratch, size_t scratch_bytes, int* counts, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	if (P < 0 || !counts) return E3_ERR_BAD_ARGUMENT;
	if (split_n != 2) return E3_ERR_UNSUPPORTED;
	if (P >= (1 << 30)) return E3_ERR_TOO_LARGE;
	if (P == 0) { DCHECK(cudaMemsetAsync(counts, 0, 8 * sizeof(int), st)); return 0; }
	if (!xyz_gradient_accum || !denom || !scaling || !opacity || !scratch || scratch_bytes < e3_densify_scratch_bytes(P))
		return E3_ERR_BAD_ARGUMENT;
	DensifyScratch s = DensifyScratch::carve(static_cast<char*>(scratch), (size_t)P);
	const float inv_split = 1.0f / (float)(0.8 * split_n);   // ATen divides by a scalar as a * (1 / b)
	const int blocks = (P + kThreads - 1) / kThreads;
	DCHECK(cudaMemsetAsync(counts, 0, 8 * sizeof(int), st));
	densify_flags_kernel<<<blocks, kThreads, 0, st>>>(P, xyz_gradient_accum, denom, scaling, opacity, max_grad, dense_extent,
		min_opacity, big_world_extent, inv_split, s);
	uint32_t* c = reinterpret_cast<uint32_t*>(counts);
	DCHECK(e3_exclusive_scan_u32(s.flag_keep, nullptr, s.off_keep, (size_t)P, s.scan_tmp, c + 0, st));
	DCHECK(e3_exclusive_scan_u32(s.flag_clone, nullptr, s.off_clone, (size_t)P, s.scan_tmp, c + 1, st));
	DCHECK(e3_exclusive_scan_u32(s.flag_child, nullptr, s.off_child, (size_t)P, s.scan_tmp, c + 2, st));
	DCHECK(e3_exclusive_scan_u32(s.flag_sel, nullptr, s.rank_sel, (size_t)P, s.scan_tmp, c + 3, st));
	densify_map_kernel<<<blocks, kThreads, 0, st>>>(P, s, counts);
	g_e3_launches += 2;
	return (int)cudaGetLastError();
}

extern "C" int e3_densify_apply(int P, int P_new, int num_split_selected, const e3_densify_array* arrays_host, int num_arrays,
	const float* rotation, const float* scaling, const float* normals, int split_n, const void* scratch, size_t scratch_bytes,
	void* stream)
{
	if (P < 0 || P_new < 0 || P_new > 2 * (long long)P || num_arrays < 0 || num_arrays > E3_DENSIFY_MAX_ARRAYS ||
		num_split_selected < 0 || num_split_selected > P)
		return E3_ERR_BAD_ARGUMENT;
	if (split_n != 2) return E3_ERR_UNSUPPORTED;
	if (P_new == 0 || num_arrays == 0) return 0;
	if (!arrays_host || !scratch || scratch_bytes < e3_densify_scratch_bytes(P)) return E3_ERR_BAD_ARGUMENT;
	DensifyScratch s = DensifyScratch::carve(const_cast<char*>(static_cast<const char*>(scratch)), (size_t)P);
	GatherTable tab;
	tab.num = 0;
	unsigned blocks = 0;
	for (int i = 0; i < num_arrays; i++) {
		const e3_densify_array& a = arrays_host[i];
		if (!a.src || !a.dst || a.width <= 0 || a.kind < 0 || a.kind > 2) return E3_ERR_BAD_ARGUMENT;
		if ((a.kind == 1 || a.kind == 2) && a.width != 3) return E3_ERR_BAD_ARGUMENT;
		if (a.kind == 1 && num_split_selected > 0 && (!rotation || !scaling || !normals)) return E3_ERR_BAD_ARGUMENT;
		const int k = tab.num++;
		tab.src[k] = a.src; tab.dst[k] = a.dst;
		tab.m_src[k] = a.exp_avg_src; tab.m_dst[k] = a.exp_avg_dst;
		tab.v_src[k] = a.exp_avg_sq_src; tab.v_dst[k] = a.exp_avg_sq_dst;
		tab.width[k] = a.width; tab.kind[k] = a.kind;
		tab.first_block[k] = blocks;
		blocks += (unsigned)(((unsigned long long)P_new * a.width + kGatherPerBlock - 1) / kGatherPerBlock);
	}
	tab.first_block[tab.num] = blocks;
	densify_gather_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(tab, (unsigned long long)P_new, s.row_map, s.rank_sel,
		rotation, scaling, normals, (unsigned)num_split_selected, 1.0f / (float)(0.8 * split_n));
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}

extern "C" int e3_densification_stats(int P, const int* radii, const float* d_mean2D, float* max_radii2D,
	float* xyz_gradient_accum, float* denom, void* stream)
{
	if (P < 0) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (!radii || !d_mean2D || !max_radii2D || !xyz_gradient_accum || !denom) return E3_ERR_BAD_ARGUMENT;
	densification_stats_kernel<<<(P + kThreads - 1) / kThreads, kThreads, 0, (cudaStream_t)stream>>>(P, radii, d_mean2D,
		max_radii2D, xyz_gradient_accum, denom);
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}

extern "C" int e3_reset_opacity(int P, float* opacity, float* exp_avg, float* exp_avg_sq, float max_opacity, void* stream)
{
	if (P < 0 || !(max_opacity > 0.0f && max_opacity < 1.0f)) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (!opacity) return E3_ERR_BAD_ARGUMENT;
	reset_opacity_kernel<<<(P + kThreads - 1) / kThreads, kThreads, 0, (cudaStream_t)stream>>>(P, opacity, exp_avg, exp_avg_sq,
		max_opacity);
	g_e3_launches += 1;
	return (int)cudaGetLastError();
}
