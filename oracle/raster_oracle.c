/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the differentiable Gaussian rasterizer hot path.
 *
 * Plain-C restatement of the reference algorithm (lanpokn/Event-3DGS, vendored
 * diff-gaussian-rasterization @ 59f5f77). Every function cites the reference file:line it
 * follows; paths are relative to submodules/diff-gaussian-rasterization/cuda_rasterizer/.
 * Nothing in the product path (event_3dgs_b200/) may import, link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Rounding: the reference is compiled by nvcc with FMA contraction on. The geometry chain that
 * feeds the bit-exact outputs (depth key bits, radii, tile rectangles) is written here with
 * explicit fmaf() in exactly the places where the reference's sm_100a SASS has FFMA (decoded
 * from `cuobjdump -sass` of the oracle/_ref build), and this file is compiled with
 * -ffp-contract=off so the C compiler adds no fusion of its own. The generic nvcc pattern is
 *     a*b + c*d + e*f   ->  fmaf(e, f, fmaf(a, b, c*d))
 * i.e. the first product is fused, the second one is rounded on its own.
 *
 * Parity pin: the reference has no tests or golden vectors. This oracle is
 * pinned against the reference CUDA build itself (oracle/_ref/libref_dgr.so) on the GPU box by
 * tests/test_ref_parity_gpu.py, and against golden vectors captured from that build
 * (tests/golden/) by the CPU suite.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLOCK_X 16
#define BLOCK_Y 16
#define BLOCK_SIZE 256

/* auxiliary.h:22-39 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
	-1.0925484305920792f, 0.5462742152960396f };
static const float SH_C3[] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
	0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };

/* CUDA float->int conversion (cvt.rzi.s32.f32) saturates and maps NaN to 0. */
static int f2i_trunc(float v)
{
	if (v != v) return 0;
	if (v >= 2147483648.0f) return 2147483647;
	if (v <= -2147483648.0f) return (-2147483647 - 1);
	return (int)v;
}
static int f2i_ceil(float v) { return f2i_trunc(ceilf(v)); }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* a*b + c*d + e*f as nvcc contracts it */
static float dot3c(float a, float b, float c, float d, float e, float f)
{
	return fmaf(e, f, fmaf(a, b, c * d));
}

/* auxiliary.h:41-44 -- note the double-precision literals: the math is done in double,
 * and (v+1)*S-1 is one DFMA in the reference SASS. */
static float ndc2Pix(float v, int S)
{
	return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}

/* auxiliary.h:46-56. grid is unsigned in the reference; min(unsigned,int) on non-negative
 * values is the plain minimum. */
static void getRect(float px, float py, int max_radius, int gridx, int gridy,
	uint32_t* rminx, uint32_t* rminy, uint32_t* rmaxx, uint32_t* rmaxy)
{
	float r = (float)max_radius;
	*rminx = (uint32_t)imin(gridx, imax(0, f2i_trunc((px - r) * 0.0625f)));
	*rminy = (uint32_t)imin(gridy, imax(0, f2i_trunc((py - r) * 0.0625f)));
	*rmaxx = (uint32_t)imin(gridx, imax(0, f2i_trunc((((px + r) + 16.0f) - 1.0f) * 0.0625f)));
	*rmaxy = (uint32_t)imin(gridy, imax(0, f2i_trunc((((py + r) + 16.0f) - 1.0f) * 0.0625f)));
}

/* auxiliary.h:58-66; row k of the (transposed-storage) matrix */
static float xform_row(const float* m, int k, float x, float y, float z)
{
	return dot3c(m[k], x, m[4 + k], y, m[8 + k], z) + m[12 + k];
}

/* forward.cu:118-152 */
static void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D)
{
	float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
	float sx = mod * scale[0], sy = mod * scale[1], sz = mod * scale[2];
	/* rotation entries, contraction exactly as in the reference SASS */
	float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
	float A = fmaf(r, y, xz);    /* x*z + r*y */
	float B = fmaf(-r, y, xz);   /* x*z - r*y */
	float Cc = fmaf(y, z, -rx);  /* y*z - r*x */
	float Dd = fmaf(y, z, rx);   /* y*z + r*x */
	float E = fmaf(x, y, -rz);   /* x*y - r*z */
	float F = fmaf(x, y, rz);    /* x*y + r*z */
	float G = fmaf(x, x, yy);    /* x*x + y*y */
	float Hh = yy + zz;          /* y*y + z*z */
	float I = fmaf(x, x, zz);    /* x*x + z*z */
	/* glm column-major R (forward.cu:134-138): R[c][r] */
	float R[3][3];
	R[0][0] = 1.0f - (Hh + Hh); R[0][1] = E + E;            R[0][2] = A + A;
	R[1][0] = F + F;            R[1][1] = 1.0f - (I + I);   R[1][2] = Cc + Cc;
	R[2][0] = B + B;            R[2][1] = Dd + Dd;          R[2][2] = 1.0f - (G + G);
	/* M = S * R (glm) : M[c][r] = s_r * R[c][r] (the zero terms of S add nothing) */
	float s[3] = { sx, sy, sz };
	float M[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) M[c][rr] = s[rr] * R[c][rr];
	/* Sigma = transpose(M) * M : Sigma[c][r] = sum_k M[r][k] * M[c][k] */
#define SIG(c, rr) dot3c(M[rr][0], M[c][0], M[rr][1], M[c][1], M[rr][2], M[c][2])
	cov3D[0] = SIG(0, 0); cov3D[1] = SIG(0, 1); cov3D[2] = SIG(0, 2);
	cov3D[3] = SIG(1, 1); cov3D[4] = SIG(1, 2); cov3D[5] = SIG(2, 2);
#undef SIG
}

/* shared by forward.cu:74-113 and backward.cu:166-199: builds T (glm T[c][r], c<2) and returns
 * the clamped t. */
static void cov2d_T(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy,
	const float* vm, float t[3], float T[2][3], float* txtz_out, float* tytz_out)
{
	t[0] = xform_row(vm, 0, mean[0], mean[1], mean[2]);
	t[1] = xform_row(vm, 1, mean[0], mean[1], mean[2]);
	t[2] = xform_row(vm, 2, mean[0], mean[1], mean[2]);
	const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
	const float txtz = t[0] / t[2], tytz = t[1] / t[2];
	t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
	t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
	if (txtz_out) *txtz_out = txtz;
	if (tytz_out) *tytz_out = tytz;
	float tz2 = t[2] * t[2];
	float a = fx / t[2], c = -(fx * t[0]) / tz2;   /* J column 0 = (a, 0, c) */
	float b = fy / t[2], d = -(fy * t[1]) / tz2;   /* J column 1 = (0, b, d) */
	/* W[k][r] = vm[4*r + k]  (forward.cu:94-97);  T = W * J */
	for (int r = 0; r < 3; r++) {
		float W0 = vm[4 * r + 0], W1 = vm[4 * r + 1], W2 = vm[4 * r + 2];
		T[0][r] = fmaf(W2, c, fmaf(W0, a, W1 * 0.0f));
		T[1][r] = fmaf(W2, d, fmaf(W0, 0.0f, W1 * b));
	}
}

/* forward.cu:74-113 */
static void computeCov2D(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy,
	const float* c3, const float* vm, float cov[3])
{
	float t[3], T[2][3];
	cov2d_T(mean, fx, fy, tan_fovx, tan_fovy, vm, t, T, 0, 0);
	float V[3][3] = { { c3[0], c3[1], c3[2] }, { c3[1], c3[3], c3[4] }, { c3[2], c3[4], c3[5] } };
	/* X = transpose(T) * transpose(Vrk): X[c][r] = sum_k T[r][k] * V[c][k], r < 2 */
	float X[3][2];
	for (int c = 0; c < 3; c++) for (int r = 0; r < 2; r++)
		X[c][r] = dot3c(T[r][0], V[c][0], T[r][1], V[c][1], T[r][2], V[c][2]);
	/* cov = X * T: cov[c][r] = sum_k X[k][r] * T[c][k] */
#define COV(c, r) dot3c(X[0][r], T[c][0], X[1][r], T[c][1], X[2][r], T[c][2])
	cov[0] = COV(0, 0) + 0.3f;
	cov[1] = COV(0, 1);
	cov[2] = COV(1, 1) + 0.3f;
#undef COV
}

/* forward.cu:20-71 */
static void computeColorFromSH(int idx, int deg, int max_coeffs, const float* means,
	const float* campos, const float* shs, unsigned char* clamped, float out[3])
{
	float dx = means[3 * idx] - campos[0], dy = means[3 * idx + 1] - campos[1], dz = means[3 * idx + 2] - campos[2];
	float len = sqrtf(fmaf(dz, dz, fmaf(dx, dx, dy * dy)));
	float x = dx / len, y = dy / len, z = dz / len;
	const float* sh = shs + (size_t)idx * max_coeffs * 3;
	for (int c = 0; c < 3; c++) {
#define S(k) sh[3 * (k) + c]
		float res = SH_C0 * S(0);
		if (deg > 0) {
			res = fmaf(-(SH_C1 * y), S(1), res);
			res = fmaf(SH_C1 * z, S(2), res);
			res = fmaf(-(SH_C1 * x), S(3), res);
			if (deg > 1) {
				float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
				res = fmaf(SH_C2[0] * xy, S(4), res);
				res = fmaf(SH_C2[1] * yz, S(5), res);
				res = fmaf(SH_C2[2] * (((zz + zz) - xx) - yy), S(6), res);
				res = fmaf(SH_C2[3] * xz, S(7), res);
				res = fmaf(SH_C2[4] * (xx - yy), S(8), res);
				if (deg > 2) {
					res = fmaf((SH_C3[0] * y) * fmaf(xx, 3.0f, -yy), S(9), res);
					res = fmaf((SH_C3[1] * xy) * z, S(10), res);
					res = fmaf((SH_C3[2] * y) * (fmaf(zz, 4.0f, -xx) - yy), S(11), res);
					res = fmaf((SH_C3[3] * z) * fmaf(yy, -3.0f, fmaf(xx, -3.0f, zz + zz)), S(12), res);
					res = fmaf((SH_C3[4] * x) * (fmaf(zz, 4.0f, -xx) - yy), S(13), res);
					res = fmaf((SH_C3[5] * z) * (xx - yy), S(14), res);
					res = fmaf((SH_C3[6] * x) * fmaf(yy, -3.0f, xx), S(15), res);
				}
			}
		}
#undef S
		res += 0.5f;
		clamped[3 * idx + c] = (res < 0);
		out[c] = res > 0.0f ? res : 0.0f;
	}
}

/* rasterizer_impl.cu:35-50 */
uint32_t orc_get_higher_msb(uint32_t n)
{
	uint32_t msb = sizeof(n) * 4;
	uint32_t step = msb;
	while (step > 1) {
		step /= 2;
		if (n >> msb) msb += step; else msb -= step;
	}
	if (n >> msb) msb++;
	return msb;
}

/* Host threads of the OpenMP loops below: set explicitly by the timing harness (an inherited
 * OMP_NUM_THREADS=1, torchrun's default, would otherwise silently serialise the CPU baseline). */
#ifdef _OPENMP
#include <omp.h>
int orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); return omp_get_max_threads(); }
#else
int orc_set_threads(int n) { (void)n; return 1; }
#endif

/* rasterizer_impl.cu:54-66 + auxiliary.h:139-164 (markVisible / checkFrustum) */
void orc_mark_visible(int P, const float* means, const float* vm, const float* pm, unsigned char* present)
{
	(void)pm;
	for (int i = 0; i < P; i++)
		present[i] = xform_row(vm, 2, means[3 * i], means[3 * i + 1], means[3 * i + 2]) > 0.2f;
}

/* forward.cu:155-256 (preprocessCUDA). Returns -1 if a `prefiltered` point was culled
 * (the reference __trap()s, auxiliary.h:156-160). */
int orc_preprocess(int P, int D, int M, const float* means, const float* scales, float mod,
	const float* rots, const float* opac, const float* shs, unsigned char* clamped,
	const float* cov3D_precomp, const float* colors_precomp, const float* vm, const float* pm,
	const float* campos, int W, int H, float tan_fovx, float tan_fovy, float focal_x,
	float focal_y, int* radii, float* means2D, float* depths, float* cov3Ds, float* rgb,
	float* conic_opacity, int gridx, int gridy, uint32_t* tiles_touched, int prefiltered)
{
	int trap = 0;
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		radii[idx] = 0;
		tiles_touched[idx] = 0;
		const float px = means[3 * idx], py = means[3 * idx + 1], pz = means[3 * idx + 2];
		/* in_frustum, auxiliary.h:139-164 */
		float pview_z = xform_row(vm, 2, px, py, pz);
		if (!(pview_z > 0.2f)) { if (prefiltered) trap = 1; continue; }
		float hx = xform_row(pm, 0, px, py, pz), hy = xform_row(pm, 1, px, py, pz), hw = xform_row(pm, 3, px, py, pz);
		float p_w = 1.0f / (hw + 0.0000001f);
		float projx = hx * p_w, projy = hy * p_w;
		const float* cov3D;
		if (cov3D_precomp) cov3D = cov3D_precomp + 6 * (size_t)idx;
		else { computeCov3D(scales + 3 * (size_t)idx, mod, rots + 4 * (size_t)idx, cov3Ds + 6 * (size_t)idx); cov3D = cov3Ds + 6 * (size_t)idx; }
		float cov[3];
		computeCov2D(means + 3 * (size_t)idx, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, vm, cov);
		float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
		if (det == 0.0f) continue;
		float det_inv = 1.f / det;
		float conic[3] = { cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv };
		float mid = 0.5f * (cov[0] + cov[2]);
		float sq = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
		float lambda1 = mid + sq, lambda2 = mid - sq;
		int my_radius = f2i_ceil(3.f * sqrtf(fmaxf(lambda1, lambda2)));
		float pix = ndc2Pix(projx, W), piy = ndc2Pix(projy, H);
		uint32_t rminx, rminy, rmaxx, rmaxy;
		getRect(pix, piy, my_radius, gridx, gridy, &rminx, &rminy, &rmaxx, &rmaxy);
		if ((rmaxx - rminx) * (rmaxy - rminy) == 0) continue;
		if (!colors_precomp) {
			float c[3];
			computeColorFromSH(idx, D, M, means, campos, shs, clamped, c);
			rgb[3 * idx] = c[0]; rgb[3 * idx + 1] = c[1]; rgb[3 * idx + 2] = c[2];
		}
		depths[idx] = pview_z;
		radii[idx] = my_radius;
		means2D[2 * idx] = pix; means2D[2 * idx + 1] = piy;
		conic_opacity[4 * idx] = conic[0]; conic_opacity[4 * idx + 1] = conic[1];
		conic_opacity[4 * idx + 2] = conic[2]; conic_opacity[4 * idx + 3] = opac[idx];
		tiles_touched[idx] = (rmaxy - rminy) * (rmaxx - rminx);
	}
	return trap ? -1 : 0;
}

/* rasterizer_impl.cu:277 (cub::DeviceScan::InclusiveSum) */
void orc_inclusive_sum(int P, const uint32_t* in, uint32_t* out)
{
	uint32_t s = 0;
	for (int i = 0; i < P; i++) { s += in[i]; out[i] = s; }
}

/* rasterizer_impl.cu:70-111 */
void orc_duplicate_with_keys(int P, const float* means2D, const float* depths, const uint32_t* offsets,
	uint64_t* keys, uint32_t* values, const int* radii, int gridx, int gridy)
{
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (radii[idx] > 0) {
			uint32_t off = (idx == 0) ? 0 : offsets[idx - 1];
			uint32_t rminx, rminy, rmaxx, rmaxy;
			getRect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gridx, gridy, &rminx, &rminy, &rmaxx, &rmaxy);
			uint32_t dbits; memcpy(&dbits, &depths[idx], 4);
			for (uint32_t y = rminy; y < rmaxy; y++) for (uint32_t x = rminx; x < rmaxx; x++) {
				uint64_t key = (uint64_t)(y * (uint32_t)gridx + x);
				key <<= 32; key |= dbits;
				keys[off] = key; values[off] = (uint32_t)idx; off++;
			}
		}
	}
}

/* rasterizer_impl.cu:300-308: stable LSD radix sort of (key,value) pairs over key bits
 * [0,end_bit) (what cub::DeviceRadixSort::SortPairs computes). tmp arrays are caller-owned. */
void orc_sort_pairs(int R, const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in,
	uint32_t* vals_out, int end_bit, uint64_t* ktmp, uint32_t* vtmp)
{
	const uint64_t* ks = keys_in; const uint32_t* vs = vals_in;
	int npass = (end_bit + 7) / 8;
	if (npass == 0 || R == 0) { memcpy(keys_out, keys_in, 8 * (size_t)R); memcpy(vals_out, vals_in, 4 * (size_t)R); return; }
	/* ping-pong so that the last pass lands in *_out */
	uint64_t* kd = (npass & 1) ? keys_out : ktmp; uint32_t* vd = (npass & 1) ? vals_out : vtmp;
	for (int p = 0; p < npass; p++) {
		int lo = 8 * p, nb = end_bit - lo < 8 ? end_bit - lo : 8;
		uint32_t mask = (1u << nb) - 1u;
		size_t cnt[257]; memset(cnt, 0, sizeof(cnt));
		for (int i = 0; i < R; i++) cnt[((ks[i] >> lo) & mask) + 1]++;
		for (int b = 0; b < 256; b++) cnt[b + 1] += cnt[b];
		for (int i = 0; i < R; i++) { size_t d = cnt[(ks[i] >> lo) & mask]++; kd[d] = ks[i]; vd[d] = vs[i]; }
		ks = kd; vs = vd;
		kd = (kd == keys_out) ? ktmp : keys_out; vd = (vd == vals_out) ? vtmp : vals_out;
	}
}

/* rasterizer_impl.cu:116-138 (+ the memset at :310, done by the caller) */
void orc_identify_tile_ranges(int L, const uint64_t* keys, uint32_t* ranges)
{
	for (int idx = 0; idx < L; idx++) {
		uint32_t currtile = (uint32_t)(keys[idx] >> 32);
		if (idx == 0) ranges[2 * currtile] = 0;
		else {
			uint32_t prevtile = (uint32_t)(keys[idx - 1] >> 32);
			if (currtile != prevtile) { ranges[2 * prevtile + 1] = idx; ranges[2 * currtile] = idx; }
		}
		if (idx == L - 1) ranges[2 * currtile + 1] = L;
	}
}

/* Timing aid for bench.py's bounded CPU sample: restrict both blend loops to the pixel window
 * [begin, end) (row-major pixel ids). Default = whole image. Not used by any parity test. */
static int g_pix_begin = 0, g_pix_end = 0x7fffffff;
void orc_set_pixel_window(int begin, int end) { g_pix_begin = begin; g_pix_end = end; }

/* forward.cu:261-374 (renderCUDA). One pixel at a time: the per-pixel result does not depend on
 * the block's batching. C = channels of `features` (3 in the reference, config.h:15; 1 is the
 * luma rewrite of SURVEY.md section 7). */
void orc_render_forward(int C, int gridx, int gridy, const uint32_t* ranges, const uint32_t* point_list,
	int W, int H, const float* means2D, const float* features, const float* conic_opacity,
	float* final_T, uint32_t* n_contrib, const float* bg, float* out_color)
{
	(void)gridy;
	const int pb = g_pix_begin > 0 ? g_pix_begin : 0, pe = g_pix_end < W * H ? g_pix_end : W * H;
#pragma omp parallel for schedule(dynamic, 64)
	for (int pix_id = pb; pix_id < pe; pix_id++) {
		int px = pix_id % W, py = pix_id / W;
		uint32_t tile = (uint32_t)(py / BLOCK_Y) * gridx + (px / BLOCK_X);
		uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
		float pixfx = (float)px, pixfy = (float)py;
		float T = 1.0f; uint32_t contributor = 0, last_contributor = 0;
		float Cc[3] = { 0, 0, 0 };
		for (uint32_t k = r0; k < r1; k++) {
			contributor++;
			uint32_t id = point_list[k];
			float dx = means2D[2 * id] - pixfx, dy = means2D[2 * id + 1] - pixfy;
			const float* co = conic_opacity + 4 * (size_t)id;
			/* forward.cu:335, contraction as in the reference SASS */
			float sum = fmaf(dx, co[0] * dx, (co[2] * dy) * dy);
			float power = fmaf(sum, -0.5f, -((co[1] * dx) * dy));
			if (power > 0.0f) continue;
			float alpha = fminf(0.99f, co[3] * expf(power));
			if (!(alpha >= 1.0f / 255.0f)) continue;
			float test_T = T * (1 - alpha);
			if (!(test_T >= 0.0001f)) break;   /* done = true (forward.cu:347-351) */
			for (int ch = 0; ch < C; ch++) Cc[ch] = fmaf(T, alpha * features[(size_t)id * C + ch], Cc[ch]);
			T = test_T;
			last_contributor = contributor;
		}
		final_T[pix_id] = T;
		n_contrib[pix_id] = last_contributor;
		for (int ch = 0; ch < C; ch++) out_color[(size_t)ch * H * W + pix_id] = fmaf(bg[ch], T, Cc[ch]);
	}
}

/* backward.cu:399-557 (renderCUDA backward). Accumulates in double (the reference uses float
 * atomics in arbitrary order, so its own result is only defined up to summation order).
 * Outputs are float arrays that the caller zero-fills (rasterize_points.cu:151-159):
 * dL_dmean2D [P,3] (x,y written), dL_dconic [P,4] (.x .y .w), dL_dopacity [P], dL_dcolors [P,C]. */
void orc_render_backward(int C, int P, int gridx, int gridy, const uint32_t* ranges,
	const uint32_t* point_list, int W, int H, const float* bg, const float* means2D,
	const float* conic_opacity, const float* colors, const float* final_Ts,
	const uint32_t* n_contrib, const float* dL_dpixels, float* dL_dmean2D, float* dL_dconic,
	float* dL_dopacity, float* dL_dcolors)
{
	(void)gridy;
	double* acc = (double*)calloc((size_t)P * (6 + C), sizeof(double));
	const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
	const int pb = g_pix_begin > 0 ? g_pix_begin : 0, pe = g_pix_end < W * H ? g_pix_end : W * H;
#pragma omp parallel for schedule(dynamic, 64)
	for (int pix_id = pb; pix_id < pe; pix_id++) {
		int px = pix_id % W, py = pix_id / W;
		uint32_t tile = (uint32_t)(py / BLOCK_Y) * gridx + (px / BLOCK_X);
		uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
		float pixfx = (float)px, pixfy = (float)py;
		const float T_final = final_Ts[pix_id];
		float T = T_final;
		uint32_t contributor = r1 - r0;
		const uint32_t last_contributor = n_contrib[pix_id];
		float accum_rec[3] = { 0, 0, 0 }, last_color[3] = { 0, 0, 0 }, dL_dpixel[3] = { 0, 0, 0 };
		for (int ch = 0; ch < C; ch++) dL_dpixel[ch] = dL_dpixels[(size_t)ch * H * W + pix_id];
		float last_alpha = 0;
		for (uint32_t k = r1; k-- > r0;) {
			contributor--;
			if (contributor >= last_contributor) continue;
			uint32_t id = point_list[k];
			float dx = means2D[2 * id] - pixfx, dy = means2D[2 * id + 1] - pixfy;
			const float* co = conic_opacity + 4 * (size_t)id;
			float sum = fmaf(dx, co[0] * dx, (co[2] * dy) * dy);
			float power = fmaf(sum, -0.5f, -((co[1] * dx) * dy));
			if (power > 0.0f) continue;
			const float G = expf(power);
			const float alpha = fminf(0.99f, co[3] * G);
			if (!(alpha >= 1.0f / 255.0f)) continue;
			T = T / (1.f - alpha);
			const float dchannel_dcolor = alpha * T;
			float dL_dalpha = 0.0f;
			double* a = acc + (size_t)id * (6 + C);
			for (int ch = 0; ch < C; ch++) {
				const float c = colors[(size_t)id * C + ch];
				accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
				last_color[ch] = c;
				const float dL_dchannel = dL_dpixel[ch];
				dL_dalpha += (c - accum_rec[ch]) * dL_dchannel;
				double v = (double)(dchannel_dcolor * dL_dchannel);
#pragma omp atomic
				a[6 + ch] += v;
			}
			dL_dalpha *= T;
			last_alpha = alpha;
			float bg_dot_dpixel = 0;
			for (int ch = 0; ch < C; ch++) bg_dot_dpixel += bg[ch] * dL_dpixel[ch];
			dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
			const float dL_dG = co[3] * dL_dalpha;
			const float gdx = G * dx, gdy = G * dy;
			const float dG_ddelx = -gdx * co[0] - gdy * co[1];
			const float dG_ddely = -gdy * co[2] - gdx * co[1];
			double v0 = (double)(dL_dG * dG_ddelx * ddelx_dx), v1 = (double)(dL_dG * dG_ddely * ddely_dy);
			double v2 = (double)(-0.5f * gdx * dx * dL_dG), v3 = (double)(-0.5f * gdx * dy * dL_dG);
			double v4 = (double)(-0.5f * gdy * dy * dL_dG), v5 = (double)(G * dL_dalpha);
#pragma omp atomic
			a[0] += v0;
#pragma omp atomic
			a[1] += v1;
#pragma omp atomic
			a[2] += v2;
#pragma omp atomic
			a[3] += v3;
#pragma omp atomic
			a[4] += v4;
#pragma omp atomic
			a[5] += v5;
		}
	}
	for (int i = 0; i < P; i++) {
		const double* a = acc + (size_t)i * (6 + C);
		dL_dmean2D[3 * i] += (float)a[0]; dL_dmean2D[3 * i + 1] += (float)a[1];
		dL_dconic[4 * i] += (float)a[2]; dL_dconic[4 * i + 1] += (float)a[3]; dL_dconic[4 * i + 3] += (float)a[4];
		dL_dopacity[i] += (float)a[5];
		for (int ch = 0; ch < C; ch++) dL_dcolors[(size_t)i * C + ch] += (float)a[6 + ch];
	}
	free(acc);
}

/* backward.cu:144-274 (computeCov2DCUDA). glm T[c][r]; the reference's T[i][j] is that. */
void orc_cov2d_backward(int P, const float* means, const int* radii, const float* cov3Ds,
	float h_x, float h_y, float tan_fovx, float tan_fovy, const float* vm, const float* dL_dconics,
	float* dL_dmeans, float* dL_dcov)
{
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (!(radii[idx] > 0)) continue;
		const float* c3 = cov3Ds + 6 * (size_t)idx;
		float dcx = dL_dconics[4 * idx], dcy = dL_dconics[4 * idx + 1], dcz = dL_dconics[4 * idx + 3];
		float t[3], T2[2][3], txtz, tytz;
		cov2d_T(means + 3 * (size_t)idx, h_x, h_y, tan_fovx, tan_fovy, vm, t, T2, &txtz, &tytz);
		const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
		const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
		const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
		float T[3][3] = { { T2[0][0], T2[0][1], T2[0][2] }, { T2[1][0], T2[1][1], T2[1][2] }, { 0, 0, 0 } };
		float V[3][3] = { { c3[0], c3[1], c3[2] }, { c3[1], c3[3], c3[4] }, { c3[2], c3[4], c3[5] } };
		float W[3][3];   /* W[k][r] = vm[4r+k], backward.cu:182-185 */
		for (int k = 0; k < 3; k++) for (int r = 0; r < 3; r++) W[k][r] = vm[4 * r + k];
		float X[3][2];
		for (int c = 0; c < 3; c++) for (int r = 0; r < 2; r++)
			X[c][r] = T[r][0] * V[c][0] + T[r][1] * V[c][1] + T[r][2] * V[c][2];
		float a = X[0][0] * T[0][0] + X[1][0] * T[0][1] + X[2][0] * T[0][2] + 0.3f;
		float b = X[0][1] * T[0][0] + X[1][1] * T[0][1] + X[2][1] * T[0][2];
		float c = X[0][1] * T[1][0] + X[1][1] * T[1][1] + X[2][1] * T[1][2] + 0.3f;
		float denom = a * c - b * b;
		float dL_da = 0, dL_db = 0, dL_dc = 0;
		float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
		float* dcov = dL_dcov + 6 * (size_t)idx;
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
		} else {
			for (int i = 0; i < 6; i++) dcov[i] = 0;
		}
		float dL_dT00 = 2 * (T[0][0] * V[0][0] + T[0][1] * V[0][1] + T[0][2] * V[0][2]) * dL_da + (T[1][0] * V[0][0] + T[1][1] * V[0][1] + T[1][2] * V[0][2]) * dL_db;
		float dL_dT01 = 2 * (T[0][0] * V[1][0] + T[0][1] * V[1][1] + T[0][2] * V[1][2]) * dL_da + (T[1][0] * V[1][0] + T[1][1] * V[1][1] + T[1][2] * V[1][2]) * dL_db;
		float dL_dT02 = 2 * (T[0][0] * V[2][0] + T[0][1] * V[2][1] + T[0][2] * V[2][2]) * dL_da + (T[1][0] * V[2][0] + T[1][1] * V[2][1] + T[1][2] * V[2][2]) * dL_db;
		float dL_dT10 = 2 * (T[1][0] * V[0][0] + T[1][1] * V[0][1] + T[1][2] * V[0][2]) * dL_dc + (T[0][0] * V[0][0] + T[0][1] * V[0][1] + T[0][2] * V[0][2]) * dL_db;
		float dL_dT11 = 2 * (T[1][0] * V[1][0] + T[1][1] * V[1][1] + T[1][2] * V[1][2]) * dL_dc + (T[0][0] * V[1][0] + T[0][1] * V[1][1] + T[0][2] * V[1][2]) * dL_db;
		float dL_dT12 = 2 * (T[1][0] * V[2][0] + T[1][1] * V[2][1] + T[1][2] * V[2][2]) * dL_dc + (T[0][0] * V[2][0] + T[0][1] * V[2][1] + T[0][2] * V[2][2]) * dL_db;
		float dL_dJ00 = W[0][0] * dL_dT00 + W[0][1] * dL_dT01 + W[0][2] * dL_dT02;
		float dL_dJ02 = W[2][0] * dL_dT00 + W[2][1] * dL_dT01 + W[2][2] * dL_dT02;
		float dL_dJ11 = W[1][0] * dL_dT10 + W[1][1] * dL_dT11 + W[1][2] * dL_dT12;
		float dL_dJ12 = W[2][0] * dL_dT10 + W[2][1] * dL_dT11 + W[2][2] * dL_dT12;
		float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
		float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
		float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
		float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t[0]) * tz3 * dL_dJ02 + (2 * h_y * t[1]) * tz3 * dL_dJ12;
		/* transformVec4x3Transpose, auxiliary.h:89-97; ASSIGNED (backward.cu:273) */
		dL_dmeans[3 * idx + 0] = vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz;
		dL_dmeans[3 * idx + 1] = vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz;
		dL_dmeans[3 * idx + 2] = vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz;
	}
}

/* backward.cu:20-139 (computeColorFromSH backward) */
static void sh_backward(int idx, int deg, int max_coeffs, const float* means, const float* campos,
	const float* shs, const unsigned char* clamped, const float* dL_dcolor, float* dL_dmeans, float* dL_dshs)
{
	float dox = means[3 * idx] - campos[0], doy = means[3 * idx + 1] - campos[1], doz = means[3 * idx + 2] - campos[2];
	float len = sqrtf(dox * dox + doy * doy + doz * doz);
	float x = dox / len, y = doy / len, z = doz / len;
	const float* sh = shs + (size_t)idx * max_coeffs * 3;
	float* dsh = dL_dshs + (size_t)idx * max_coeffs * 3;
	float dRGB[3];
	for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[3 * idx + c] * (clamped[3 * idx + c] ? 0.f : 1.f);
	float dRGBdx[3] = { 0, 0, 0 }, dRGBdy[3] = { 0, 0, 0 }, dRGBdz[3] = { 0, 0, 0 };
#define S(k) sh[3 * (k) + c]
#define DSH(k, w) for (int c = 0; c < 3; c++) dsh[3 * (k) + c] = (w) * dRGB[c]
	DSH(0, SH_C0);
	if (deg > 0) {
		DSH(1, -SH_C1 * y); DSH(2, SH_C1 * z); DSH(3, -SH_C1 * x);
		for (int c = 0; c < 3; c++) { dRGBdx[c] = -SH_C1 * S(3); dRGBdy[c] = -SH_C1 * S(1); dRGBdz[c] = SH_C1 * S(2); }
		if (deg > 1) {
			float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
			DSH(4, SH_C2[0] * xy); DSH(5, SH_C2[1] * yz); DSH(6, SH_C2[2] * (2.f * zz - xx - yy));
			DSH(7, SH_C2[3] * xz); DSH(8, SH_C2[4] * (xx - yy));
			for (int c = 0; c < 3; c++) {
				dRGBdx[c] += SH_C2[0] * y * S(4) + SH_C2[2] * 2.f * -x * S(6) + SH_C2[3] * z * S(7) + SH_C2[4] * 2.f * x * S(8);
				dRGBdy[c] += SH_C2[0] * x * S(4) + SH_C2[1] * z * S(5) + SH_C2[2] * 2.f * -y * S(6) + SH_C2[4] * 2.f * -y * S(8);
				dRGBdz[c] += SH_C2[1] * y * S(5) + SH_C2[2] * 2.f * 2.f * z * S(6) + SH_C2[3] * x * S(7);
			}
			if (deg > 2) {
				DSH(9, SH_C3[0] * y * (3.f * xx - yy)); DSH(10, SH_C3[1] * xy * z);
				DSH(11, SH_C3[2] * y * (4.f * zz - xx - yy)); DSH(12, SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy));
				DSH(13, SH_C3[4] * x * (4.f * zz - xx - yy)); DSH(14, SH_C3[5] * z * (xx - yy));
				DSH(15, SH_C3[6] * x * (xx - 3.f * yy));
				for (int c = 0; c < 3; c++) {
					dRGBdx[c] += (SH_C3[0] * S(9) * 3.f * 2.f * xy + SH_C3[1] * S(10) * yz + SH_C3[2] * S(11) * -2.f * xy +
						SH_C3[3] * S(12) * -3.f * 2.f * xz + SH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
						SH_C3[5] * S(14) * 2.f * xz + SH_C3[6] * S(15) * 3.f * (xx - yy));
					dRGBdy[c] += (SH_C3[0] * S(9) * 3.f * (xx - yy) + SH_C3[1] * S(10) * xz + SH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) +
						SH_C3[3] * S(12) * -3.f * 2.f * yz + SH_C3[4] * S(13) * -2.f * xy + SH_C3[5] * S(14) * -2.f * yz +
						SH_C3[6] * S(15) * -3.f * 2.f * xy);
					dRGBdz[c] += (SH_C3[1] * S(10) * xy + SH_C3[2] * S(11) * 4.f * 2.f * yz + SH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) +
						SH_C3[4] * S(13) * 4.f * 2.f * xz + SH_C3[5] * S(14) * (xx - yy));
				}
			}
		}
	}
#undef S
#undef DSH
	float ddx = dRGBdx[0] * dRGB[0] + dRGBdx[1] * dRGB[1] + dRGBdx[2] * dRGB[2];
	float ddy = dRGBdy[0] * dRGB[0] + dRGBdy[1] * dRGB[1] + dRGBdy[2] * dRGB[2];
	float ddz = dRGBdz[0] * dRGB[0] + dRGBdz[1] * dRGB[1] + dRGBdz[2] * dRGB[2];
	/* dnormvdv, auxiliary.h:107-117 */
	float sum2 = dox * dox + doy * doy + doz * doz;
	float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
	dL_dmeans[3 * idx + 0] += ((+sum2 - dox * dox) * ddx - doy * dox * ddy - doz * dox * ddz) * invsum32;
	dL_dmeans[3 * idx + 1] += (-dox * doy * ddx + (sum2 - doy * doy) * ddy - doz * doy * ddz) * invsum32;
	dL_dmeans[3 * idx + 2] += (-dox * doz * ddx - doy * doz * ddy + (sum2 - doz * doz) * ddz) * invsum32;
}

/* backward.cu:278-341 (computeCov3D backward) */
static void cov3d_backward(int idx, const float* scale, float mod, const float* rot, const float* dL_dcov3Ds,
	float* dL_dscales, float* dL_drots)
{
	float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
	/* glm R[c][r] */
	float R[3][3] = {
		{ 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
		{ 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
		{ 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
	float s[3] = { mod * scale[0], mod * scale[1], mod * scale[2] };
	float M[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) M[c][rr] = s[rr] * R[c][rr];
	const float* d = dL_dcov3Ds + 6 * (size_t)idx;
	float dS[3][3] = { { d[0], 0.5f * d[1], 0.5f * d[2] }, { 0.5f * d[1], d[3], 0.5f * d[4] }, { 0.5f * d[2], 0.5f * d[4], d[5] } };
	/* dL_dM = 2.0f * M * dL_dSigma  (glm: (2*M) * dS) : dM[c][rr] = sum_k 2M[k][rr] * dS[c][k] */
	float dM[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++)
		dM[c][rr] = (2.0f * M[0][rr]) * dS[c][0] + (2.0f * M[1][rr]) * dS[c][1] + (2.0f * M[2][rr]) * dS[c][2];
	/* Rt = transpose(R), dL_dMt = transpose(dL_dM); Rt[i] . dL_dMt[i] = sum_k R[k][i] * dM[k][i] */
	float dMt[3][3];
	for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) dMt[c][rr] = dM[rr][c];
	for (int i = 0; i < 3; i++)
		dL_dscales[3 * idx + i] = R[0][i] * dMt[i][0] + R[1][i] * dMt[i][1] + R[2][i] * dMt[i][2];
	for (int i = 0; i < 3; i++) for (int k = 0; k < 3; k++) dMt[i][k] *= s[i];
	float* q = dL_drots + 4 * (size_t)idx;
	q[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
	q[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
	q[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
	q[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
}

/* backward.cu:346-396 (preprocessCUDA backward). shs / scales may be NULL. */
void orc_preprocess_backward(int P, int D, int M, const float* means, const int* radii, const float* shs,
	const unsigned char* clamped, const float* scales, const float* rots, float mod, const float* proj,
	const float* campos, const float* dL_dmean2D, float* dL_dmeans, const float* dL_dcolor,
	const float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot)
{
#pragma omp parallel for schedule(static)
	for (int idx = 0; idx < P; idx++) {
		if (!(radii[idx] > 0)) continue;
		float mx = means[3 * idx], my = means[3 * idx + 1], mz = means[3 * idx + 2];
		float m_homw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
		float m_w = 1.0f / (m_homw + 0.0000001f);
		float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
		float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
		float gx = dL_dmean2D[3 * idx], gy = dL_dmean2D[3 * idx + 1];
		dL_dmeans[3 * idx + 0] += (proj[0] * m_w - proj[3] * mul1) * gx + (proj[1] * m_w - proj[3] * mul2) * gy;
		dL_dmeans[3 * idx + 1] += (proj[4] * m_w - proj[7] * mul1) * gx + (proj[5] * m_w - proj[7] * mul2) * gy;
		dL_dmeans[3 * idx + 2] += (proj[8] * m_w - proj[11] * mul1) * gx + (proj[9] * m_w - proj[11] * mul2) * gy;
		if (shs) sh_backward(idx, D, M, means, campos, shs, clamped, dL_dcolor, dL_dmeans, dL_dsh);
		if (scales) cov3d_backward(idx, scales + 3 * (size_t)idx, mod, rots + 4 * (size_t)idx, dL_dcov3D, dL_dscale, dL_drot);
	}
}
