// Per-Gaussian stages of the splatting path: projection / covariance / extent (forward) and
// the gradient chain back to the leaf attributes (backward). sm_100a.
//
// Forward follows the arithmetic of the reference preprocessCUDA
// (cuda_rasterizer/forward.cu:155-256 with computeCov3D :118-152, computeCov2D :74-113 and the
// helpers of auxiliary.h:41-77,139-164) ROUNDING FOR ROUNDING: radii, tile rectangles and depth
// key bits have to be bit-identical to the reference build, so every multiply-add is written
// with an explicit __fmaf_rn / __fmul_rn / __fadd_rn in the contraction pattern that nvcc emits
// for the reference source (decoded from its sm_100a SASS). The generic pattern for
// a*b + c*d + e*f is fma(e,f, fma(a,b, rn(c*d))).
#include "common.cuh"

namespace {

__device__ __forceinline__ float dot3c(float a, float b, float c, float d, float e, float f)
{
	return __fmaf_rn(e, f, __fmaf_rn(a, b, __fmul_rn(c, d)));
}

// row k of a matrix stored transposed (auxiliary.h:58-77)
__device__ __forceinline__ float xform_row(const float* __restrict__ m, int k, float x, float y, float z)
{
	return __fadd_rn(dot3c(m[k], x, m[4 + k], y, m[8 + k], z), m[12 + k]);
}

// auxiliary.h:41-44 (double arithmetic, one DFMA)
__device__ __forceinline__ float ndc2pix(float v, int S)
{
	return (float)__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5);
}

// auxiliary.h:46-56; float->int conversions saturate like the reference's cvt.rzi
__device__ __forceinline__ void get_rect(float px, float py, int max_radius, int gx, int gy,
	uint32_t& x0, uint32_t& y0, uint32_t& x1, uint32_t& y1)
{
	const float r = (float)max_radius;
	x0 = min(gx, max(0, __float2int_rz(__fmul_rn(__fadd_rn(px, -r), 0.0625f))));
	y0 = min(gy, max(0, __float2int_rz(__fmul_rn(__fadd_rn(py, -r), 0.0625f))));
	x1 = min(gx, max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, r), 16.0f), -1.0f), 0.0625f))));
	y1 = min(gy, max(0, __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, r), 16.0f), -1.0f), 0.0625f))));
}

// forward.cu:118-152
__device__ __forceinline__ void cov3d_exact(float sx, float sy, float sz, float mod, float r, float x,
	float y, float z, float* c3)
{
	sx = __fmul_rn(mod, sx); sy = __fmul_rn(mod, sy); sz = __fmul_rn(mod, sz);
	const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
	const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
	const float A = __fmaf_rn(r, y, xz), B = __fmaf_rn(-r, y, xz);
	const float Cc = __fmaf_rn(y, z, -rx), Dd = __fmaf_rn(y, z, rx);
	const float E = __fmaf_rn(x, y, -rz), F = __fmaf_rn(x, y, rz);
	const float G = __fmaf_rn(x, x, yy), Hh = __fadd_rn(yy, zz), I = __fmaf_rn(x, x, zz);
	// glm column-major R[c][r], then M[c][r] = s_r * R[c][r]
	const float M00 = __fmul_rn(sx, __fadd_rn(1.0f, -__fadd_rn(Hh, Hh)));
	const float M01 = __fmul_rn(sy, __fadd_rn(E, E));
	const float M02 = __fmul_rn(sz, __fadd_rn(A, A));
	const float M10 = __fmul_rn(sx, __fadd_rn(F, F));
	const float M11 = __fmul_rn(sy, __fadd_rn(1.0f, -__fadd_rn(I, I)));
	const float M12 = __fmul_rn(sz, __fadd_rn(Cc, Cc));
	const float M20 = __fmul_rn(sx, __fadd_rn(B, B));
	const float M21 = __fmul_rn(sy, __fadd_rn(Dd, Dd));
	const float M22 = __fmul_rn(sz, __fadd_rn(1.0f, -__fadd_rn(G, G)));
	// Sigma[c][r] = sum_k M[r][k] * M[c][k]
	c3[0] = dot3c(M00, M00, M01, M01, M02, M02);
	c3[1] = dot3c(M10, M00, M11, M01, M12, M02);
	c3[2] = dot3c(M20, M00, M21, M01, M22, M02);
	c3[3] = dot3c(M10, M10, M11, M11, M12, M12);
	c3[4] = dot3c(M20, M10, M21, M11, M22, M12);
	c3[5] = dot3c(M20, M20, M21, M21, M22, M22);
}

// forward.cu:74-113
__device__ __forceinline__ float3 cov2d_exact(float mx, float my, float mz, float fx, float fy,
	float tan_fovx, float tan_fovy, const float* c3, const float* __restrict__ vm)
{
	float tx = xform_row(vm, 0, mx, my, mz), ty = xform_row(vm, 1, mx, my, mz);
	const float tz = xform_row(vm, 2, mx, my, mz);
	const float limx = __fmul_rn(1.3f, tan_fovx), limy = __fmul_rn(1.3f, tan_fovy);
	const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
	tx = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), tz);
	ty = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), tz);
	const float tz2 = __fmul_rn(tz, tz);
	const float a = __fdiv_rn(fx, tz), c = __fdiv_rn(-__fmul_rn(fx, tx), tz2);
	const float b = __fdiv_rn(fy, tz), d = __fdiv_rn(-__fmul_rn(fy, ty), tz2);
	float T0[3], T1[3];   // glm T[0][r], T[1][r]
#pragma unroll
	for (int r = 0; r < 3; r++) {
		const float W0 = vm[4 * r], W1 = vm[4 * r + 1], W2 = vm[4 * r + 2];
		T0[r] = __fmaf_rn(W2, c, __fmaf_rn(W0, a, __fmul_rn(W1, 0.0f)));
		T1[r] = __fmaf_rn(W2, d, __fmaf_rn(W0, 0.0f, __fmul_rn(W1, b)));
	}
	const float V[3][3] = { { c3[0], c3[1], c3[2] }, { c3[1], c3[3], c3[4] }, { c3[2], c3[4], c3[5] } };
	float X0[3], X1[3];   // X[c][0], X[c][1]
#pragma unroll
	for (int cc = 0; cc < 3; cc++) {
		X0[cc] = dot3c(T0[0], V[cc][0], T0[1], V[cc][1], T0[2], V[cc][2]);
		X1[cc] = dot3c(T1[0], V[cc][0], T1[1], V[cc][1], T1[2], V[cc][2]);
	}
	float3 cov;
	cov.x = __fadd_rn(dot3c(X0[0], T0[0], X0[1], T0[1], X0[2], T0[2]), 0.3f);
	cov.y = dot3c(X1[0], T0[0], X1[1], T0[1], X1[2], T0[2]);
	cov.z = __fadd_rn(dot3c(X1[0], T1[0], X1[1], T1[1], X1[2], T1[2]), 0.3f);
	return cov;
}

__constant__ float kSH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
	-1.0925484305920792f, 0.5462742152960396f };
__constant__ float kSH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
	0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };
#define E3_LUMA_R 0.4124f
#define E3_LUMA_G 0.35758f
#define E3_LUMA_B 0.1804f
#define SH_C0 0.28209479177387814f
#define SH_C1 0.4886025119029199f

// forward.cu:20-71
// sh0 = the degree-0 coefficient [3], shr = coefficients 1.. [(M-1),3] (contiguous after sh0 in
// the reference's [P,M,3] layout; a separate tensor, features_rest, in the raw-parameter path)
__device__ __forceinline__ float3 sh_to_color(int deg, const float* __restrict__ sh0, const float* __restrict__ shr,
	float dx, float dy, float dz, uint8_t* clamped)
{
	const float len = sqrtf(dx * dx + dy * dy + dz * dz);
	const float x = dx / len, y = dy / len, z = dz / len;
	float res[3];
#pragma unroll
	for (int c = 0; c < 3; c++) {
#define S(k) ((k) == 0 ? sh0[c] : shr[3 * ((k) - 1) + c])
		float v = SH_C0 * S(0);
		if (deg > 0) {
			v = v - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
			if (deg > 1) {
				const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
				v = v + kSH_C2[0] * xy * S(4) + kSH_C2[1] * yz * S(5) + kSH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
					kSH_C2[3] * xz * S(7) + kSH_C2[4] * (xx - yy) * S(8);
				if (deg > 2) {
					v = v + kSH_C3[0] * y * (3.0f * xx - yy) * S(9) + kSH_C3[1] * xy * z * S(10) +
						kSH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
						kSH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
						kSH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + kSH_C3[5] * z * (xx - yy) * S(14) +
						kSH_C3[6] * x * (xx - 3.0f * yy) * S(15);
				}
			}
		}
#undef S
		v += 0.5f;
		clamped[c] = (v < 0);
		res[c] = fmaxf(v, 0.0f);
	}
	return make_float3(res[0], res[1], res[2]);
}

// Stage `n` consecutive floats of a [P,K] AoS array for this block through shared memory with
// 16-byte loads (K = 3: a 12-byte row cannot be vector-loaded per thread).
template <int K>
__device__ __forceinline__ void stage_rows(const float* __restrict__ src, size_t first_row, int rows, float* sm)
{
	const float* base = src + first_row * K;
	const int nfl = rows * K;
	if ((reinterpret_cast<size_t>(base) & 15) == 0) {
		const int nv = nfl >> 2;
		for (int i = threadIdx.x; i < nv; i += blockDim.x)
			reinterpret_cast<float4*>(sm)[i] = __ldg(reinterpret_cast<const float4*>(base) + i);
		for (int i = (nv << 2) + threadIdx.x; i < nfl; i += blockDim.x) sm[i] = __ldg(base + i);
	} else {
		for (int i = threadIdx.x; i < nfl; i += blockDim.x) sm[i] = __ldg(base + i);
	}
}

// RAW = the inputs are the leaf parameters of scene/gaussian_model.py (log-scales, unnormalised
// quaternions, opacity logits, features_dc / features_rest): the activations of
// gaussian_model.py:95-115 and the SH -> colour of gaussian_renderer/__init__.py:77-81 are
// evaluated here instead of by ~60 PyTorch launches per render (K15 of SURVEY.md section 2b).
template <int C, bool RAW>
__global__ void __launch_bounds__(256) preprocess_fwd_kernel(
	int P, int D, int M,
	const float* __restrict__ means3D, const float* __restrict__ scales, float mod,
	const float* __restrict__ rotations, const float* __restrict__ opacities,
	const float* __restrict__ shs, const float* __restrict__ f_rest, const float* __restrict__ cov3D_precomp,
	const float* __restrict__ colors_precomp, const float* __restrict__ vm_g,
	const float* __restrict__ pm_g, const float* __restrict__ campos,
	int W, int H, float tan_fovx, float tan_fovy, float focal_x, float focal_y, int gx, int gy,
	bool prefiltered, int* __restrict__ radii, float* __restrict__ depths, float4* __restrict__ rec,
	float* __restrict__ cov3Ds, float* __restrict__ rgb, uint8_t* __restrict__ clamped,
	uint32_t* __restrict__ tiles_touched, uint2* __restrict__ rect, uint32_t* __restrict__ dkey,
	uint32_t* __restrict__ dval, float* __restrict__ acc, int* __restrict__ flags, const float* __restrict__ bg3)
{
	__shared__ float s_xyz[256 * 3];
	__shared__ float s_scl[256 * 3];
	__shared__ float s_col[256 * 3];
	__shared__ float s_m[32];
	extern __shared__ __align__(16) float s_shr[];      // [256*45] features_rest tile (RAW, degree > 0)
	const size_t first = (size_t)blockIdx.x * 256;
	const int rows = min(256, P - (int)first);
	const int t = threadIdx.x;
	// all per-thread global loads are issued before the barrier (one memory round trip)
	float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
	float opac_in = 0.f;
	if (t < rows) {
		if (!cov3D_precomp) q_in = __ldg(reinterpret_cast<const float4*>(rotations) + first + t);
		opac_in = __ldg(opacities + first + t);
	}
	stage_rows<3>(means3D, first, rows, s_xyz);
	if (scales) stage_rows<3>(scales, first, rows, s_scl);
	if (colors_precomp) stage_rows<C>(colors_precomp, first, rows, s_col);
	const bool sh_tile = RAW && D > 0;
	if (sh_tile) stage_rows<45>(f_rest, first, rows, s_shr);
	if (threadIdx.x < 16) s_m[threadIdx.x] = vm_g[threadIdx.x];
	else if (threadIdx.x < 32) s_m[threadIdx.x] = pm_g[threadIdx.x - 16];
	__syncthreads();
	if (t >= rows) return;
	const size_t idx = first + t;
	const float* vm = s_m;
	const float* pm = s_m + 16;
	if (RAW && C == 1 && idx == 0)      // luma of the 3-channel background, kept with the state for both blend passes
		reinterpret_cast<float*>(flags)[2] = E3_LUMA_R * bg3[0] + E3_LUMA_G * bg3[1] + E3_LUMA_B * bg3[2];
	if (RAW) {    // scaling_activation = exp (gaussian_model.py:28)
		s_scl[3 * t] = expf(s_scl[3 * t]); s_scl[3 * t + 1] = expf(s_scl[3 * t + 1]); s_scl[3 * t + 2] = expf(s_scl[3 * t + 2]);
	}

	// zero this Gaussian's backward accumulator while we are streaming over it anyway
	float4* a4 = reinterpret_cast<float4*>(acc + idx * E3_ACC_STRIDE);
	a4[0] = a4[1] = a4[2] = make_float4(0.f, 0.f, 0.f, 0.f);

	int my_radius = 0;
	uint32_t tt = 0, key = 0xffffffffu;
	uint2 rc = make_uint2(0, 0);
	const float px = s_xyz[3 * t], py = s_xyz[3 * t + 1], pz = s_xyz[3 * t + 2];
	const float pview_z = xform_row(vm, 2, px, py, pz);
	do {
		if (!(pview_z > 0.2f)) {     // in_frustum, auxiliary.h:154
			if (prefiltered) atomicExch(&flags[1], 1);
			break;
		}
		const float hx = xform_row(pm, 0, px, py, pz), hy = xform_row(pm, 1, px, py, pz);
		const float hw = xform_row(pm, 3, px, py, pz);
		const float p_w = __frcp_rn(__fadd_rn(hw, 0.0000001f));
		const float projx = __fmul_rn(hx, p_w), projy = __fmul_rn(hy, p_w);
		float c3[6];
		if (cov3D_precomp) {
#pragma unroll
			for (int i = 0; i < 6; i++) c3[i] = cov3D_precomp[idx * 6 + i];
		} else {
			float4 q = q_in;
			if (RAW) {    // rotation_activation = F.normalize (gaussian_model.py:36): x / max(|x|, 1e-12)
				const float inv = 1.0f / fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
				q.x *= inv; q.y *= inv; q.z *= inv; q.w *= inv;
			}
			cov3d_exact(s_scl[3 * t], s_scl[3 * t + 1], s_scl[3 * t + 2], mod, q.x, q.y, q.z, q.w, c3);
#pragma unroll
			for (int i = 0; i < 6; i++) cov3Ds[idx * 6 + i] = c3[i];
		}
		const float3 cov = cov2d_exact(px, py, pz, focal_x, focal_y, tan_fovx, tan_fovy, c3, vm);
		const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
		if (det == 0.0f) break;
		const float det_inv = __frcp_rn(det);
		const float3 conic = make_float3(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv), __fmul_rn(cov.x, det_inv));
		const float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
		const float sq = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
		const float lambda1 = __fadd_rn(mid, sq), lambda2 = __fadd_rn(mid, -sq);
		const int rad = __float2int_ru(__fmul_rn(3.f, __fsqrt_rn(fmaxf(lambda1, lambda2))));
		const float pix = ndc2pix(projx, W), piy = ndc2pix(projy, H);
		uint32_t x0, y0, x1, y1;
		get_rect(pix, piy, rad, gx, gy, x0, y0, x1, y1);
		if ((x1 - x0) * (y1 - y0) == 0) break;
		float col[3] = { 0.f, 0.f, 0.f };
		if (colors_precomp) {
#pragma unroll
			for (int c = 0; c < C; c++) col[c] = s_col[C * t + c];
		} else {
			uint8_t cl[3];
			const float* sh0 = RAW ? shs + idx * 3 : shs + idx * M * 3;
			const float* shr = RAW ? (sh_tile ? s_shr + 45 * t : f_rest + idx * 45) : sh0 + 3;
			const float3 v = sh_to_color(D, sh0, shr, px - campos[0], py - campos[1], pz - campos[2], cl);
			col[0] = v.x; col[1] = v.y; col[2] = v.z;
			rgb[idx * 3] = v.x; rgb[idx * 3 + 1] = v.y; rgb[idx * 3 + 2] = v.z;
			clamped[idx * 3] = cl[0]; clamped[idx * 3 + 1] = cl[1]; clamped[idx * 3 + 2] = cl[2];
			// single-channel mode of the raw path: blend the luma of utils/loss_utils.py:25 directly
			// (blending is linear in colour, so this equals the luma of the RGB render)
			if (RAW && C == 1) col[0] = E3_LUMA_R * v.x + E3_LUMA_G * v.y + E3_LUMA_B * v.z;
		}
		my_radius = rad;
		tt = (y1 - y0) * (x1 - x0);
		rc = make_uint2(x0 | (x1 << 16), y0 | (y1 << 16));
		key = __float_as_uint(pview_z);
		depths[idx] = pview_z;
		rec[3 * idx] = make_float4(pix, piy, conic.x, conic.y);
		// half-extent of the bounding box of { alpha >= 1/255 } = { power >= -ln(255 opacity) }:
		// sqrt(2 tau cov_xx), sqrt(2 tau cov_yy) (cov = inverse of the conic). Margins make the
		// box a strict superset under fp32 rounding of the blend's own power evaluation.
		float opac = opac_in;
		if (RAW) opac = 1.0f / (1.0f + expf(-opac));      // opacity_activation = sigmoid
		const float tau = __logf(255.0f * opac) + 1e-3f;
		float ext_x = -1.f, ext_y = -1.f;
		if (tau > 0.f) {
			ext_x = sqrtf(2.0f * tau * cov.x) * 1.0005f + 0.02f;
			ext_y = sqrtf(2.0f * tau * cov.z) * 1.0005f + 0.02f;
		}
		rec[3 * idx + 1] = make_float4(conic.z, opac, col[0], col[1]);
		rec[3 * idx + 2] = make_float4(col[2], ext_x, ext_y, tau);
		if (RAW) {
			// Fused path only (its instance lists are not observable): the reference's rect comes
			// from the 3-sigma radius, but a pixel is touched only inside the { alpha >= 1/255 } box
			// above, which for low opacities is much smaller. Tiles outside it would be sorted,
			// loaded and then skipped by every pixel, so they are not emitted at all (about 28 % of
			// the instances of the BASELINE c3 scene). Results do not change by a bit; radii (the
			// visibility filter) stay the reference's.
			if (!(ext_x >= 0.f)) {
				tt = 0;
			} else {
				const int tx0 = max((int)x0, (int)floorf((pix - ext_x) * (1.0f / E3_TILE_X)));
				const int tx1 = min((int)x1, (int)floorf((pix + ext_x) * (1.0f / E3_TILE_X)) + 1);
				const int ty0 = max((int)y0, (int)floorf((piy - ext_y) * (1.0f / E3_TILE_Y)));
				const int ty1 = min((int)y1, (int)floorf((piy + ext_y) * (1.0f / E3_TILE_Y)) + 1);
				if (tx1 <= tx0 || ty1 <= ty0) {
					tt = 0;
				} else {
					tt = (uint32_t)((ty1 - ty0) * (tx1 - tx0));
					rc = make_uint2((uint32_t)tx0 | ((uint32_t)tx1 << 16), (uint32_t)ty0 | ((uint32_t)ty1 << 16));
				}
			}
		}
	} while (0);
	radii[idx] = my_radius;
	tiles_touched[idx] = tt;
	rect[idx] = rc;
	dkey[idx] = key;
	dval[idx] = (uint32_t)idx;
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ vm,
	uint8_t* __restrict__ present)
{
	const int idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= P) return;
	present[idx] = xform_row(vm, 2, means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]) > 0.2f;
}

// ---------------------------------------------------------------------------------------------
// Backward: one kernel for what the reference does in computeCov2DCUDA (backward.cu:144-274)
// followed by preprocessCUDA (backward.cu:346-396, with the SH and cov3D backwards of :20-139 and
// :278-341). Reads the packed accumulators of the backward blend and writes EVERY output row
// (zeros for culled Gaussians), so the caller does not pre-zero anything.
// ---------------------------------------------------------------------------------------------
// dsh0 / dshr: where the gradients of coefficient 0 / coefficients 1.. go. ACC adds to what is
// there (the raw path sums the three renders of an event iteration into one gradient tensor).
template <bool ACC>
__device__ __forceinline__ void sh_backward(int deg, int M, const float* __restrict__ sh0, const float* __restrict__ shr,
	float dox, float doy, float doz, const uint8_t* __restrict__ cl, const float* dL_dcolor, float* __restrict__ dsh0,
	float* __restrict__ dshr, float3& dmean)
{
	const float len = sqrtf(dox * dox + doy * doy + doz * doz);
	const float x = dox / len, y = doy / len, z = doz / len;
	float dRGB[3];
#pragma unroll
	for (int c = 0; c < 3; c++) dRGB[c] = cl[c] ? 0.f : dL_dcolor[c];
	float dRGBdx[3] = { 0, 0, 0 }, dRGBdy[3] = { 0, 0, 0 }, dRGBdz[3] = { 0, 0, 0 };
#define S(k) ((k) == 0 ? sh0[c] : shr[3 * ((k) - 1) + c])
#define DSH(k, w) { const float ww = (w); float* dd = (k) == 0 ? dsh0 : dshr + 3 * ((k) - 1); \
	_Pragma("unroll") for (int c = 0; c < 3; c++) { if (ACC) dd[c] += ww * dRGB[c]; else dd[c] = ww * dRGB[c]; } }
	DSH(0, SH_C0);
	int written = 1;
	if (deg > 0) {
		DSH(1, -SH_C1 * y); DSH(2, SH_C1 * z); DSH(3, -SH_C1 * x);
		written = 4;
#pragma unroll
		for (int c = 0; c < 3; c++) { dRGBdx[c] = -SH_C1 * S(3); dRGBdy[c] = -SH_C1 * S(1); dRGBdz[c] = SH_C1 * S(2); }
		if (deg > 1) {
			const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			DSH(4, kSH_C2[0] * xy); DSH(5, kSH_C2[1] * yz); DSH(6, kSH_C2[2] * (2.f * zz - xx - yy));
			DSH(7, kSH_C2[3] * xz); DSH(8, kSH_C2[4] * (xx - yy));
			written = 9;
#pragma unroll
			for (int c = 0; c < 3; c++) {
				dRGBdx[c] += kSH_C2[0] * y * S(4) + kSH_C2[2] * 2.f * -x * S(6) + kSH_C2[3] * z * S(7) + kSH_C2[4] * 2.f * x * S(8);
				dRGBdy[c] += kSH_C2[0] * x * S(4) + kSH_C2[1] * z * S(5) + kSH_C2[2] * 2.f * -y * S(6) + kSH_C2[4] * 2.f * -y * S(8);
				dRGBdz[c] += kSH_C2[1] * y * S(5) + kSH_C2[2] * 2.f * 2.f * z * S(6) + kSH_C2[3] * x * S(7);
			}
			if (deg > 2) {
				DSH(9, kSH_C3[0] * y * (3.f * xx - yy)); DSH(10, kSH_C3[1] * xy * z);
				DSH(11, kSH_C3[2] * y * (4.f * zz - xx - yy)); DSH(12, kSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
				DSH(13, kSH_C3[4] * x * (4.f * zz - xx - yy)); DSH(14, kSH_C3[5] * z * (xx - yy));
				DSH(15, kSH_C3[6] * x * (xx - 3.f * yy));
				written = 16;
#pragma unroll
				for (int c = 0; c < 3; c++) {
					dRGBdx[c] += (kSH_C3[0] * S(9) * 3.f * 2.f * xy + kSH_C3[1] * S(10) * yz + kSH_C3[2] * S(11) * -2.f * xy +
						kSH_C3[3] * S(12) * -3.f * 2.f * xz + kSH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
						kSH_C3[5] * S(14) * 2.f * xz + kSH_C3[6] * S(15) * 3.f * (xx - yy));
					dRGBdy[c] += (kSH_C3[0] * S(9) * 3.f * (xx - yy) + kSH_C3[1] * S(10) * xz + kSH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
						kSH_C3[3] * S(12) * -3.f * 2.f * yz + kSH_C3[4] * S(13) * -2.f * xy + kSH_C3[5] * S(14) * -2.f * yz +
						kSH_C3[6] * S(15) * -3.f * 2.f * xy);
					dRGBdz[c] += (kSH_C3[1] * S(10) * xy + kSH_C3[2] * S(11) * 4.f * 2.f * yz + kSH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
						kSH_C3[4] * S(13) * 4.f * 2.f * xz + kSH_C3[5] * S(14) * (xx - yy));
				}
			}
		}
	}
#undef S
#undef DSH
	if (!ACC && dshr)
		for (int k = written; k < M; k++) { dshr[3 * (k - 1)] = 0.f; dshr[3 * (k - 1) + 1] = 0.f; dshr[3 * (k - 1) + 2] = 0.f; }
	const float ddx = dRGBdx[0] * dRGB[0] + dRGBdx[1] * dRGB[1] + dRGBdx[2] * dRGB[2];
	const float ddy = dRGBdy[0] * dRGB[0] + dRGBdy[1] * dRGB[1] + dRGBdy[2] * dRGB[2];
	const float ddz = dRGBdz[0] * dRGB[0] + dRGBdz[1] * dRGB[1] + dRGBdz[2] * dRGB[2];
	// dnormvdv, auxiliary.h:107-117
	const float sum2 = dox * dox + doy * doy + doz * doz;
	const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
	dmean.x += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
	dmean.y += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
	dmean.z += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
}

// Accumulating writes. mc == 0: reductions into this GPU's memory (REDG, no read round trip; safe
// against the other views' kernels accumulating into the same rows concurrently). mc != 0: the same
// reduction issued on the NVLS multicast alias of the address (multimem.red): the NVSwitch applies it
// to every rank's copy of the gradient buffer, i.e. the gradient all-reduce happens in this kernel.
__device__ __forceinline__ void red_add1(float* p, float v, long long mc)
{
	if (mc) asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" :: "l"(reinterpret_cast<char*>(p) + mc), "f"(v) : "memory");
	else atomicAdd(p, v);
}
__device__ __forceinline__ void red_add4(float4* p, const float4 v, long long mc)
{
	if (mc)
		asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(reinterpret_cast<char*>(p) + mc), "f"(v.x),
			"f"(v.y), "f"(v.z), "f"(v.w) : "memory");
	else
		asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <bool ACC>
__device__ __forceinline__ void put(float* p, float v, long long mc) { if (ACC) red_add1(p, v, mc); else *p = v; }

// coalesced write-out of a staged [rows,K] tile (float4 where the global row start is aligned).
// ACC: atomic accumulation (all-zero vectors, e.g. the rows of culled Gaussians, are skipped).
template <bool ACC>
__device__ __forceinline__ void flush_rows(float* __restrict__ dst, const float* sm, int nfl, long long mc = 0)
{
	if ((reinterpret_cast<size_t>(dst) & 15) == 0) {
		const int nv = nfl >> 2;
		float4* d4 = reinterpret_cast<float4*>(dst);
		const float4* s4 = reinterpret_cast<const float4*>(sm);
		for (int i = threadIdx.x; i < nv; i += blockDim.x) {
			const float4 v = s4[i];
			if (ACC) { if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) red_add4(d4 + i, v, mc); }
			else d4[i] = v;
		}
		for (int i = (nv << 2) + threadIdx.x; i < nfl; i += blockDim.x) { if (ACC) { if (sm[i] != 0.f) red_add1(dst + i, sm[i], mc); } else dst[i] = sm[i]; }
	} else {
		for (int i = threadIdx.x; i < nfl; i += blockDim.x) { if (ACC) { if (sm[i] != 0.f) red_add1(dst + i, sm[i], mc); } else dst[i] = sm[i]; }
	}
}

#define E3_PB_THREADS 128

// RAW: gradients are taken all the way to the leaf parameters (chain rule through exp /
// normalize / sigmoid and the SH evaluation, i.e. the autograd backward of K15 in SURVEY.md 2b).
// ACC: add to the output rows (atomically) instead of overwriting them: the three renders of an event
// iteration accumulate into the same gradient tensors, from three streams. SHREST: features_rest and its
// gradient are staged through shared memory as [128,45] tiles (RAW with degree > 0).
// Pointer roles in RAW mode: shs = features_dc [P,3], f_rest = features_rest [P,15,3],
// dL_dsh = d features_dc, d_frest = d features_rest, dL_dopacity = d opacity logit,
// dL_dscale = d log-scale, dL_drot = d raw quaternion, dL_dmean2D optional,
// dL_dcolor / dL_dcov3D / dL_dconic unused (NULL).
// All 12-byte-row inputs and outputs go through shared memory so that global traffic is
// 16-byte vectorised and fully coalesced (a thread-per-row access pattern has stride 12/180 B).
template <int C, bool RAW, bool ACC, bool SHREST>
// 10 blocks of 128 threads per SM (48 registers, 48 bytes of spills) instead of the 7 that 71 registers allow: the kernel is
// bound by memory latency (44 % of the HBM rate at 40 % occupancy); 0.252 -> 0.229 ms per c3 iteration. 8: 0.241, 12: 0.238, 14: 0.285.
__global__ void __launch_bounds__(E3_PB_THREADS, 10) preprocess_bwd_kernel(
	int P, int D, int M,
	const float* __restrict__ means3D, const int* __restrict__ radii, const float* __restrict__ shs,
	const float* __restrict__ f_rest, float* __restrict__ d_frest, const float* __restrict__ opacities_raw,
	const uint8_t* __restrict__ clamped, const float* __restrict__ scales,
	const float* __restrict__ rotations, float mod, const float* __restrict__ cov3D_precomp,
	const float* __restrict__ cov3D_state, const float* __restrict__ vm_g,
	const float* __restrict__ pm_g, const float* __restrict__ campos, float h_x, float h_y,
	float tan_fovx, float tan_fovy, float* __restrict__ acc,
	float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
	float* __restrict__ dL_dcolor, float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D,
	float* __restrict__ dL_dsh, float* __restrict__ dL_dscale, float* __restrict__ dL_drot, long long mc, bool mean2d_second,
	const float4* __restrict__ rec, int W, int H)
{
	constexpr int BT = E3_PB_THREADS;
	extern __shared__ __align__(16) float smem[];
	float* s_mean = smem;               // in: means3D rows   -> out: d means3D
	float* s_scl = s_mean + BT * 3;     // in: scale rows     -> out: d scale
	float* s_fdc = s_scl + BT * 3;      // in: features_dc    -> out: d features_dc (RAW)
	float* s_cov = s_fdc + BT * 3;      // in: cov3D rows [BT*6] -> out: d mean2D [BT*3]
	float* s_m = s_cov + BT * 6;        // view + proj matrices [32]
	float* s_sh = s_m + 32;             // in: features_rest tile [BT*45] (SHREST)
	float* s_dsh = s_sh + BT * 45;      // out: d features_rest tile [BT*45] (SHREST)
	const size_t first = (size_t)blockIdx.x * BT;
	const int rows = min(BT, P - (int)first);
	const int t = threadIdx.x;
	const bool live = t < rows;
	const size_t idx = first + (live ? t : 0);
	// Issue EVERY global load of this thread before the first barrier, whether or not the
	// Gaussian turns out to be visible: one round trip to memory instead of a chain of
	// dependent ones (the kernel is latency bound, not bandwidth bound).
	int rad = 0;
	float4 a0 = make_float4(0, 0, 0, 0), a1 = a0, a2 = a0, rc1 = a0;
	float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
	float opac_raw = 0.f;
	uint8_t cl[3] = { 0, 0, 0 };
	float4* a4 = reinterpret_cast<float4*>(acc + idx * E3_ACC_STRIDE);
	if (live) {
		rad = radii[idx];
		a0 = a4[0]; a1 = a4[1]; a2 = a4[2];
		rc1 = __ldg(rec + 3 * idx + 1);
		if (scales) q = __ldg(reinterpret_cast<const float4*>(rotations) + idx);
		if (RAW) opac_raw = __ldg(opacities_raw + idx);
		if (shs) { cl[0] = clamped[3 * idx]; cl[1] = clamped[3 * idx + 1]; cl[2] = clamped[3 * idx + 2]; }
	}
	stage_rows<3>(means3D, first, rows, s_mean);
	if (scales) stage_rows<3>(scales, first, rows, s_scl);
	if (RAW) stage_rows<3>(shs, first, rows, s_fdc);
	stage_rows<6>(cov3D_precomp ? cov3D_precomp : cov3D_state, first, rows, s_cov);
	if (SHREST) stage_rows<45>(f_rest, first, rows, s_sh);
	if (threadIdx.x < 16) s_m[threadIdx.x] = vm_g[threadIdx.x];
	else if (threadIdx.x < 32) s_m[threadIdx.x] = pm_g[threadIdx.x - 16];
	__syncthreads();
	const float* vm = s_m;
	const float* proj = s_m + 16;
	const bool vis = live && rad > 0;
	if (vis) {
		// leave the accumulator clean so that a second backward of the same forward is correct
		a4[0] = a4[1] = a4[2] = make_float4(0.f, 0.f, 0.f, 0.f);
	} else {
		a0 = a1 = a2 = make_float4(0.f, 0.f, 0.f, 0.f);
	}
	if (vis) {      // (the record of a culled Gaussian was never written)
		// The backward blend leaves sums that lack their per-Gaussian factors (g = G dL/dalpha, d = centre - pixel):
		//   a0 = {sum g, -sum g dpower/ddx, -sum g dpower/ddy, sum g dx^2}  a1 = {sum g dx dy, sum g dy^2, dL/dc0, dL/dc1}
		//   a2 = {dL/dc2, second seed's -sum g dpower/ddx, -sum g dpower/ddy, -}
		// With dL/dG = opacity g this is backward.cu:538-554: dL/dmean2D = dL/dG dG/ddel (0.5 W, 0.5 H),
		// dL/dconic = -0.5 G d d^T dL/dG, dL/dopacity = sum g.
		// From here on a0 = {dL/dmean2D.x, .y, dL/dconic.xx, .xy}, a1 = {dL/dconic.yy, dL/dopacity, ..}, a2.yz = second seed's dL/dmean2D.
		const float o = rc1.y;
		const float sw = -0.5f * (float)W * o, sh = -0.5f * (float)H * o, sc = -0.5f * o;
		const float sum_g = a0.x;
		a0 = make_float4(sw * a0.y, sh * a0.z, sc * a0.w, sc * a1.x);
		a1 = make_float4(sc * a1.y, sum_g, a1.z, a1.w);
		a2.y *= sw; a2.z *= sh;
	}
	const float3 mean = make_float3(s_mean[3 * t], s_mean[3 * t + 1], s_mean[3 * t + 2]);
	float c3[6];
#pragma unroll
	for (int i = 0; i < 6; i++) c3[i] = s_cov[6 * t + i];
	float sact[3] = { 1.f, 1.f, 1.f };
	if (scales) { sact[0] = s_scl[3 * t]; sact[1] = s_scl[3 * t + 1]; sact[2] = s_scl[3 * t + 2]; }
	float fdc[3] = { 0.f, 0.f, 0.f };
	if (RAW) { fdc[0] = s_fdc[3 * t]; fdc[1] = s_fdc[3 * t + 1]; fdc[2] = s_fdc[3 * t + 2]; }
	__syncthreads();          // every staged input row is in registers: the buffers now carry outputs

	// single-channel raw mode: the blend produced dL/dY; dL/dcolour_k = w_k dL/dY
	const float dcol[3] = { (RAW && C == 1) ? E3_LUMA_R * a1.z : a1.z, (RAW && C == 1) ? E3_LUMA_G * a1.z : a1.w,
		(RAW && C == 1) ? E3_LUMA_B * a1.z : a2.x };
	if (live) {
		if (dL_dconic) reinterpret_cast<float4*>(dL_dconic)[idx] = make_float4(a0.z, a0.w, 0.f, a1.x);
		if (!RAW) {
			dL_dopacity[idx] = a1.y;
#pragma unroll
			for (int c = 0; c < C; c++) dL_dcolor[idx * C + c] = dcol[c];
		} else {
			// sigmoid backward (gaussian_model.py:31): d logit = d opacity * o * (1 - o)
			float dlogit = 0.f;
			if (vis) { const float o = 1.0f / (1.0f + expf(-opac_raw)); dlogit = a1.y * o * (1.0f - o); }
			if (!ACC || vis) put<ACC>(dL_dopacity + idx, dlogit, mc);
		}
	}
	// d mean2D row handed to the caller (densification statistics): slots 0, 1, or the second seed's 9, 10
	s_cov[3 * t] = mean2d_second ? a2.y : a0.x; s_cov[3 * t + 1] = mean2d_second ? a2.z : a0.y; s_cov[3 * t + 2] = 0.f;

	float3 dmean = make_float3(0, 0, 0);
	float dcov[6] = { 0, 0, 0, 0, 0, 0 };
	float dfdc[3] = { 0.f, 0.f, 0.f };
	if (SHREST) {
		float* row = s_dsh + 45 * t;
		for (int k = 0; k < 45; k++) row[k] = 0.f;
	}
	if (vis) {
		const float dcx = a0.z, dcy = a0.w, dcz = a1.x;
		// ---- computeCov2DCUDA, backward.cu:166-273 ----
		float3 tt = make_float3(vm[0] * mean.x + vm[4] * mean.y + vm[8] * mean.z + vm[12],
			vm[1] * mean.x + vm[5] * mean.y + vm[9] * mean.z + vm[13],
			vm[2] * mean.x + vm[6] * mean.y + vm[10] * mean.z + vm[14]);
		const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
		const float txtz = tt.x / tt.z, tytz = tt.y / tt.z;
		tt.x = fminf(limx, fmaxf(-limx, txtz)) * tt.z;
		tt.y = fminf(limy, fmaxf(-limy, tytz)) * tt.z;
		const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
		const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
		const float ja = h_x / tt.z, jc = -(h_x * tt.x) / (tt.z * tt.z);
		const float jb = h_y / tt.z, jd = -(h_y * tt.y) / (tt.z * tt.z);
		float T[2][3], Wm[3][3];    // glm T[c][r], W[k][r] = vm[4r+k]
#pragma unroll
		for (int r = 0; r < 3; r++) {
			Wm[0][r] = vm[4 * r]; Wm[1][r] = vm[4 * r + 1]; Wm[2][r] = vm[4 * r + 2];
			T[0][r] = Wm[0][r] * ja + Wm[2][r] * jc;
			T[1][r] = Wm[1][r] * jb + Wm[2][r] * jd;
		}
		const float V[3][3] = { { c3[0], c3[1], c3[2] }, { c3[1], c3[3], c3[4] }, { c3[2], c3[4], c3[5] } };
		float TV0[3], TV1[3];     // TV0[c] = sum_k T[0][k] V[c][k]
#pragma unroll
		for (int c = 0; c < 3; c++) {
			TV0[c] = T[0][0] * V[c][0] + T[0][1] * V[c][1] + T[0][2] * V[c][2];
			TV1[c] = T[1][0] * V[c][0] + T[1][1] * V[c][1] + T[1][2] * V[c][2];
		}
		const float a = TV0[0] * T[0][0] + TV0[1] * T[0][1] + TV0[2] * T[0][2] + 0.3f;
		const float b = TV1[0] * T[0][0] + TV1[1] * T[0][1] + TV1[2] * T[0][2];
		const float c = TV1[0] * T[1][0] + TV1[1] * T[1][1] + TV1[2] * T[1][2] + 0.3f;
		const float denom = a * c - b * b;
		float dL_da = 0, dL_db = 0, dL_dc = 0;
		const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
		if (denom2inv != 0) {
			dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
			dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
			dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
			dcov[0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
			dcov[3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
			dcov[5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
			dcov[1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
			dcov[2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
			dcov[4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;
		}
		// dL/dT (backward.cu:237-248): dT0c = 2*TV0[c]*dL_da + TV1[c]*dL_db, dT1c = 2*TV1[c]*dL_dc + TV0[c]*dL_db
		float dT0[3], dT1[3];
#pragma unroll
		for (int cc = 0; cc < 3; cc++) {
			dT0[cc] = 2 * TV0[cc] * dL_da + TV1[cc] * dL_db;
			dT1[cc] = 2 * TV1[cc] * dL_dc + TV0[cc] * dL_db;
		}
		const float dL_dJ00 = Wm[0][0] * dT0[0] + Wm[0][1] * dT0[1] + Wm[0][2] * dT0[2];
		const float dL_dJ02 = Wm[2][0] * dT0[0] + Wm[2][1] * dT0[1] + Wm[2][2] * dT0[2];
		const float dL_dJ11 = Wm[1][0] * dT1[0] + Wm[1][1] * dT1[1] + Wm[1][2] * dT1[2];
		const float dL_dJ12 = Wm[2][0] * dT1[0] + Wm[2][1] * dT1[1] + Wm[2][2] * dT1[2];
		const float tz = 1.f / tt.z, tz2 = tz * tz, tz3 = tz2 * tz;
		const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
		const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
		const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * tt.x) * tz3 * dL_dJ02 + (2 * h_y * tt.y) * tz3 * dL_dJ12;
		dmean.x = vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz;
		dmean.y = vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz;
		dmean.z = vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz;
		// ---- preprocessCUDA backward, backward.cu:370-387 ----
		const float m_homw = proj[3] * mean.x + proj[7] * mean.y + proj[11] * mean.z + proj[15];
		const float m_w = 1.0f / (m_homw + 0.0000001f);
		const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
		const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
		const float gx2 = a0.x, gy2 = a0.y;
		dmean.x += (proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2;
		dmean.y += (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2;
		dmean.z += (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2;
		if (shs) {
			if (RAW)
				sh_backward<false>(D, 16, fdc, SHREST ? s_sh + 45 * t : (f_rest ? f_rest + idx * 45 : nullptr), mean.x - campos[0],
					mean.y - campos[1], mean.z - campos[2], cl, dcol, dfdc,
					SHREST ? s_dsh + 45 * t : nullptr, dmean);
			else
				sh_backward<false>(D, M, shs + idx * M * 3, shs + idx * M * 3 + 3, mean.x - campos[0], mean.y - campos[1],
					mean.z - campos[2], cl, dcol, dL_dsh + idx * M * 3, dL_dsh + idx * M * 3 + 3, dmean);
		}
	} else if (live && shs && !RAW) {
		for (int k = 0; k < 3 * M; k++) dL_dsh[idx * M * 3 + k] = 0.f;
	}
	s_mean[3 * t] = dmean.x; s_mean[3 * t + 1] = dmean.y; s_mean[3 * t + 2] = dmean.z;
	if (RAW) { s_fdc[3 * t] = dfdc[0]; s_fdc[3 * t + 1] = dfdc[1]; s_fdc[3 * t + 2] = dfdc[2]; }
	if (live && dL_dcov3D) {
#pragma unroll
		for (int i = 0; i < 6; i++) dL_dcov3D[6 * idx + i] = dcov[i];
	}

	// ---- computeCov3D backward, backward.cu:278-341 ----
	if (scales) {
		float ds[3] = { 0, 0, 0 };
		float4 dq = make_float4(0, 0, 0, 0);
		float qnorm_inv = 1.f;
		if (vis) {
			if (RAW) {
				qnorm_inv = 1.0f / fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
				q.x *= qnorm_inv; q.y *= qnorm_inv; q.z *= qnorm_inv; q.w *= qnorm_inv;
				sact[0] = expf(sact[0]); sact[1] = expf(sact[1]); sact[2] = expf(sact[2]);
			}
			const float r = q.x, x = q.y, y = q.z, z = q.w;
			const float R[3][3] = {
				{ 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
				{ 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
				{ 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
			const float s[3] = { mod * sact[0], mod * sact[1], mod * sact[2] };
			float Mx[3][3];
#pragma unroll
			for (int c = 0; c < 3; c++)
#pragma unroll
				for (int rr = 0; rr < 3; rr++) Mx[c][rr] = s[rr] * R[c][rr];
			const float dS[3][3] = { { dcov[0], 0.5f * dcov[1], 0.5f * dcov[2] }, { 0.5f * dcov[1], dcov[3], 0.5f * dcov[4] },
				{ 0.5f * dcov[2], 0.5f * dcov[4], dcov[5] } };
			float dMt[3][3];   // dMt[rr][c] = dM[c][rr], dM = (2M) * dS
#pragma unroll
			for (int c = 0; c < 3; c++)
#pragma unroll
				for (int rr = 0; rr < 3; rr++)
					dMt[rr][c] = (2.0f * Mx[0][rr]) * dS[c][0] + (2.0f * Mx[1][rr]) * dS[c][1] + (2.0f * Mx[2][rr]) * dS[c][2];
#pragma unroll
			for (int i = 0; i < 3; i++) ds[i] = R[0][i] * dMt[i][0] + R[1][i] * dMt[i][1] + R[2][i] * dMt[i][2];
#pragma unroll
			for (int i = 0; i < 3; i++)
#pragma unroll
				for (int k = 0; k < 3; k++) dMt[i][k] *= s[i];
			dq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
			dq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
			dq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
			dq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
			if (RAW) {
				// exp backward: d log-scale = d scale * scale;  F.normalize backward:
				// d raw = (d qhat - qhat (qhat . d qhat)) / |raw|
#pragma unroll
				for (int i = 0; i < 3; i++) ds[i] *= sact[i];
				const float dot = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;
				dq.x = (dq.x - q.x * dot) * qnorm_inv; dq.y = (dq.y - q.y * dot) * qnorm_inv;
				dq.z = (dq.z - q.z * dot) * qnorm_inv; dq.w = (dq.w - q.w * dot) * qnorm_inv;
			}
		}
		s_scl[3 * t] = ds[0]; s_scl[3 * t + 1] = ds[1]; s_scl[3 * t + 2] = ds[2];
		if (live && (!ACC || vis)) {
			float4* dr = reinterpret_cast<float4*>(dL_drot) + idx;
			if (ACC) red_add4(dr, dq, mc); else *dr = dq;
		}
	}
	__syncthreads();
	flush_rows<ACC>(dL_dmean3D + first * 3, s_mean, rows * 3, mc);
	if (scales) flush_rows<ACC>(dL_dscale + first * 3, s_scl, rows * 3, mc);
	if (RAW) flush_rows<ACC>(dL_dsh + first * 3, s_fdc, rows * 3, mc);
	if (dL_dmean2D) flush_rows<false>(dL_dmean2D + first * 3, s_cov, rows * 3);   // this view's own screen-space gradient: assigned
	if (SHREST) flush_rows<ACC>(d_frest + first * 45, s_dsh, rows * 45, mc);
}

}  // namespace

cudaError_t e3_launch_preprocess(const E3Frame& f, const float* means3D, const float* scales,
	const float* rotations, const float* opacities, const float* shs, const float* f_rest, bool raw,
	const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix, const float* projmatrix,
	const float* cam_pos, bool prefiltered, int* radii, E3GeomState& g, const float* bg3, cudaStream_t st)
{
	const int grid = (f.P + 255) / 256;
	const size_t dyn = (raw && f.D > 0) ? (size_t)256 * 45 * sizeof(float) : 0;
	if (dyn) {
		cudaError_t ea = e3_allow_dynamic_smem(preprocess_fwd_kernel<3, true>, dyn);
		if (ea != cudaSuccess) return ea;
		ea = e3_allow_dynamic_smem(preprocess_fwd_kernel<1, true>, dyn);
		if (ea != cudaSuccess) return ea;
	}
#define LAUNCH(CH, RW) preprocess_fwd_kernel<CH, RW><<<grid, 256, dyn, st>>>(f.P, f.D, f.M, means3D, scales, f.scale_modifier, \
	rotations, opacities, shs, f_rest, cov3D_precomp, colors_precomp, viewmatrix, projmatrix, cam_pos, f.W, f.H, \
	f.tan_fovx, f.tan_fovy, f.focal_x, f.focal_y, f.gx, f.gy, prefiltered, radii, g.depths, g.rec, g.cov3D, g.rgb, \
	g.clamped, g.tiles_touched, g.rect, g.dkey[0], g.dval[0], g.acc, g.flags, bg3)
	if (raw) { if (f.C == 1) LAUNCH(1, true); else LAUNCH(3, true); }
	else if (f.C == 1) LAUNCH(1, false); else LAUNCH(3, false);
#undef LAUNCH
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_mark_visible(int P, const float* means3D, const float* viewmatrix,
	uint8_t* present, cudaStream_t st)
{
	mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, viewmatrix, present);
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_preprocess_backward(const E3Frame& f, const float* means3D,
	const int* radii, const float* shs, const float* f_rest, const float* opacities_raw, bool raw, bool accumulate,
	const float* scales, const float* rotations,
	const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
	const float* campos, const E3GeomState& g, float* dL_dmean2D, float* dL_dconic,
	float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* d_frest,
	float* dL_dscale, float* dL_drot, cudaStream_t st)
{
	constexpr int BT = E3_PB_THREADS;
	const int grid = (f.P + BT - 1) / BT;
	const bool shrest = raw && f.D > 0 && d_frest != nullptr;
	const size_t smem_small = (size_t)(BT * 15 + 32) * sizeof(float);
	const size_t smem_big = smem_small + (size_t)BT * 90 * sizeof(float);
#define LAUNCH(CH, RW, AC, SR) do { \
	auto kern = preprocess_bwd_kernel<CH, RW, AC, SR>; \
	const size_t sm = SR ? smem_big : smem_small; \
	if (SR) { cudaError_t ea = e3_allow_dynamic_smem(kern, sm); if (ea != cudaSuccess) return ea; } \
	kern<<<grid, BT, sm, st>>>(f.P, f.D, f.M, means3D, radii, shs, f_rest, d_frest, \
	opacities_raw, g.clamped, scales, rotations, f.scale_modifier, cov3D_precomp, g.cov3D, viewmatrix, projmatrix, campos, f.focal_x, \
	f.focal_y, f.tan_fovx, f.tan_fovy, g.acc, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh, \
	dL_dscale, dL_drot, (AC) ? f.mc_delta : 0ll, f.mean2d_from_second_seed, g.rec, f.W, f.H); } while (0)
	if (raw && f.C == 1) {
		if (shrest) { if (accumulate) LAUNCH(1, true, true, true); else LAUNCH(1, true, false, true); }
		else { if (accumulate) LAUNCH(1, true, true, false); else LAUNCH(1, true, false, false); }
	} else if (raw) {
		if (shrest) { if (accumulate) LAUNCH(3, true, true, true); else LAUNCH(3, true, false, true); }
		else { if (accumulate) LAUNCH(3, true, true, false); else LAUNCH(3, true, false, false); }
	}
	else if (f.C == 1) LAUNCH(1, false, false, false); else LAUNCH(3, false, false, false);
#undef LAUNCH
	g_e3_launches += 1;
	return cudaGetLastError();
}
