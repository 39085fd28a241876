// L1 + SSIM loss of the non-event training branches, value and backward seed. sm_100a.
//
// Reference: the `--gray` branch (train.py:213-223: l1_loss_gray + ssim_gray on the luma
// 0.299 R + 0.587 G + 0.114 B, utils/loss_utils.py:18-23,40-48,368-385) and the plain RGB branch
// (train.py:292-296: l1_loss + ssim, utils/loss_utils.py:270-271,388-418), both
//     loss = (1 - lambda) * L1 + lambda * (1 - SSIM),   SSIM = mean of the 11x11 Gaussian-window
// map (sigma 1.5, zero padding 5, C1 = 0.01^2, C2 = 0.03^2). There it is 5 depthwise cuDNN
// convolutions + ~15 element-wise launches and their autograd backward. Here two tiled kernels:
//   pass 1: per 16x16 tile (+5 halo) the five windowed moments by a separable 11-tap filter in
//           shared memory, the SSIM map, its sums, and the three partial derivatives
//           dS/dmu1, dS/dE[x^2], dS/dE[xy] (the window is linear in x, x^2, xy);
//   pass 2: the same filter over those three maps gives dL/dx; plus the L1 sign term; for the
//           gray branch the luma weights carry dL/dx back to the three channels.
// Algorithmic bytes per pixel and channel: pass 1 reads 8 and writes 12, pass 2 reads 20 and
// writes 4 (gray: 12 for the three channels). Shared-memory bound stencils.
#include "common.cuh"

namespace {

struct Win11 { float w[11]; };

#define SSIM_T 16
#define SSIM_H (SSIM_T + 10)
#define SSIM_C1 0.0001f
#define SSIM_C2 0.0009f

template <typename GT> __device__ __forceinline__ float gtl(const GT* __restrict__ p, size_t i);
template <> __device__ __forceinline__ float gtl<float>(const float* __restrict__ p, size_t i) { return __ldg(p + i); }
template <> __device__ __forceinline__ float gtl<uint8_t>(const uint8_t* __restrict__ p, size_t i) { return (float)__ldg(p + i) / 255.0f; }

// value of channel `ch` of the (optionally luma-projected) image at pixel (x,y); 0 outside
template <bool GRAY>
__device__ __forceinline__ float img_at(const float* __restrict__ img, int W, int H, int ch, int x, int y)
{
	if (x < 0 || y < 0 || x >= W || y >= H) return 0.f;
	const size_t N = (size_t)W * H, i = (size_t)y * W + x;
	if (GRAY) return 0.299f * __ldg(img + i) + 0.587f * __ldg(img + N + i) + 0.114f * __ldg(img + 2 * N + i);
	return __ldg(img + ch * N + i);
}
template <bool GRAY, typename GT>
__device__ __forceinline__ float gt_at(const GT* __restrict__ img, int W, int H, int ch, int x, int y)
{
	if (x < 0 || y < 0 || x >= W || y >= H) return 0.f;
	const size_t N = (size_t)W * H, i = (size_t)y * W + x;
	if (GRAY) return 0.299f * gtl<GT>(img, i) + 0.587f * gtl<GT>(img, N + i) + 0.114f * gtl<GT>(img, 2 * N + i);
	return gtl<GT>(img, ch * N + i);
}

__device__ __forceinline__ double blk_sum(double v, double* sm)
{
	for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(E3_FULL, v, o);
	const int tid = threadIdx.y * SSIM_T + threadIdx.x;
	__syncthreads();
	if ((tid & 31) == 0) sm[tid >> 5] = v;
	__syncthreads();
	double r = 0.0;
	if (tid == 0) for (int i = 0; i < 8; i++) r += sm[i];
	return r;
}

// grid (tiles_x, tiles_y, channels); block 16x16. maps: [3][channels][N]
template <bool GRAY, typename GT>
__global__ void __launch_bounds__(256) ssim_fwd_kernel(int W, int H, const float* __restrict__ img, const GT* __restrict__ gt,
	Win11 win, float* __restrict__ maps, double* __restrict__ sums)
{
	__shared__ float sx[SSIM_H][SSIM_H + 1], sy[SSIM_H][SSIM_H + 1];
	__shared__ float hz[5][SSIM_H][SSIM_T + 1];
	__shared__ double sred[8];
	const int ch = blockIdx.z, nch = gridDim.z;
	const int x0 = blockIdx.x * SSIM_T, y0 = blockIdx.y * SSIM_T;
	const int tid = threadIdx.y * SSIM_T + threadIdx.x;
	for (int i = tid; i < SSIM_H * SSIM_H; i += 256) {
		const int ly = i / SSIM_H, lx = i - ly * SSIM_H;
		sx[ly][lx] = img_at<GRAY>(img, W, H, ch, x0 + lx - 5, y0 + ly - 5);
		sy[ly][lx] = gt_at<GRAY, GT>(gt, W, H, ch, x0 + lx - 5, y0 + ly - 5);
	}
	__syncthreads();
	// horizontal pass: 26 rows x 16 columns
	for (int i = tid; i < SSIM_H * SSIM_T; i += 256) {
		const int ly = i / SSIM_T, lx = i - ly * SSIM_T;
		float a = 0, b = 0, c = 0, d = 0, e = 0;
#pragma unroll
		for (int k = 0; k < 11; k++) {
			const float w = win.w[k], x = sx[ly][lx + k], y = sy[ly][lx + k];
			a += w * x; b += w * y; c += w * x * x; d += w * y * y; e += w * x * y;
		}
		hz[0][ly][lx] = a; hz[1][ly][lx] = b; hz[2][ly][lx] = c; hz[3][ly][lx] = d; hz[4][ly][lx] = e;
	}
	__syncthreads();
	const int tx = threadIdx.x, ty = threadIdx.y;
	float mu1 = 0, mu2 = 0, exx = 0, eyy = 0, exy = 0;
#pragma unroll
	for (int k = 0; k < 11; k++) {
		const float w = win.w[k];
		mu1 += w * hz[0][ty + k][tx]; mu2 += w * hz[1][ty + k][tx]; exx += w * hz[2][ty + k][tx];
		eyy += w * hz[3][ty + k][tx]; exy += w * hz[4][ty + k][tx];
	}
	const int px = x0 + tx, py = y0 + ty;
	const bool inside = px < W && py < H;
	double s_ssim = 0.0, s_l1 = 0.0;
	if (inside) {
		const float s1 = exx - mu1 * mu1, s2 = eyy - mu2 * mu2, s12 = exy - mu1 * mu2;
		const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
		const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = s1 + s2 + SSIM_C2;
		const float inv = 1.f / (B1 * B2);
		const float S = A1 * A2 * inv;
		// partial derivatives w.r.t. the raw windowed moments mu1, E[x^2], E[xy]
		const float dmu1 = (2.f * mu2 * (A2 - A1)) * inv - S * (2.f * mu1 / B1 - 2.f * mu1 / B2);
		const float dexx = -S / B2;
		const float dexy = 2.f * A1 * inv;
		const size_t N = (size_t)W * H, i = (size_t)py * W + px;
		maps[(0 * nch + ch) * N + i] = dmu1;
		maps[(1 * nch + ch) * N + i] = dexx;
		maps[(2 * nch + ch) * N + i] = dexy;
		s_ssim = S;
		s_l1 = fabsf(sx[ty + 5][tx + 5] - sy[ty + 5][tx + 5]);
	}
	s_ssim = blk_sum(s_ssim, sred);
	s_l1 = blk_sum(s_l1, sred);
	if (tid == 0) { atomicAdd(sums + 0, s_ssim); atomicAdd(sums + 1, s_l1); }
}

__device__ __forceinline__ float map_at(const float* __restrict__ m, int W, int H, int x, int y)
{
	return (x < 0 || y < 0 || x >= W || y >= H) ? 0.f : __ldg(m + (size_t)y * W + x);
}

// out3 = { loss, L1, SSIM }
template <bool GRAY, typename GT>
__global__ void __launch_bounds__(256) ssim_bwd_kernel(int W, int H, const float* __restrict__ img, const GT* __restrict__ gt,
	Win11 win, const float* __restrict__ maps, const double* __restrict__ sums, float lambda, float upstream,
	float* __restrict__ d_image, float* __restrict__ out3)
{
	__shared__ float sm[3][SSIM_H][SSIM_H + 1];
	__shared__ float hz[3][SSIM_H][SSIM_T + 1];
	const int ch = blockIdx.z, nch = gridDim.z;
	const size_t N = (size_t)W * H;
	const int x0 = blockIdx.x * SSIM_T, y0 = blockIdx.y * SSIM_T;
	const int tid = threadIdx.y * SSIM_T + threadIdx.x;
	if (blockIdx.x == 0 && blockIdx.y == 0 && ch == 0 && tid == 0) {
		const double cnt = (double)N * nch;
		const double ssim = sums[0] / cnt, l1 = sums[1] / cnt;
		out3[0] = (float)((1.0 - lambda) * l1 + lambda * (1.0 - ssim));
		out3[1] = (float)l1; out3[2] = (float)ssim;
	}
	for (int i = tid; i < SSIM_H * SSIM_H; i += 256) {
		const int ly = i / SSIM_H, lx = i - ly * SSIM_H;
#pragma unroll
		for (int m = 0; m < 3; m++) sm[m][ly][lx] = map_at(maps + (m * nch + ch) * N, W, H, x0 + lx - 5, y0 + ly - 5);
	}
	__syncthreads();
	for (int i = tid; i < SSIM_H * SSIM_T; i += 256) {
		const int ly = i / SSIM_T, lx = i - ly * SSIM_T;
		float a = 0, b = 0, c = 0;
#pragma unroll
		for (int k = 0; k < 11; k++) { const float w = win.w[k]; a += w * sm[0][ly][lx + k]; b += w * sm[1][ly][lx + k]; c += w * sm[2][ly][lx + k]; }
		hz[0][ly][lx] = a; hz[1][ly][lx] = b; hz[2][ly][lx] = c;
	}
	__syncthreads();
	const int tx = threadIdx.x, ty = threadIdx.y, px = x0 + tx, py = y0 + ty;
	if (px >= W || py >= H) return;
	float c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
	for (int k = 0; k < 11; k++) { const float w = win.w[k]; c1 += w * hz[0][ty + k][tx]; c2 += w * hz[1][ty + k][tx]; c3 += w * hz[2][ty + k][tx]; }
	const float x = img_at<GRAY>(img, W, H, ch, px, py), y = gt_at<GRAY, GT>(gt, W, H, ch, px, py);
	const float cnt = (float)N * nch;
	const float ks = -upstream * lambda / cnt;              // d loss / d S(q)
	const float kl = upstream * (1.f - lambda) / cnt;       // d loss / d |x - y|
	const float g = ks * (c1 + 2.f * x * c2 + y * c3) + kl * (float)((x > y) - (x < y));
	const size_t i = (size_t)py * W + px;
	if (GRAY) { d_image[i] = 0.299f * g; d_image[N + i] = 0.587f * g; d_image[2 * N + i] = 0.114f * g; }
	else d_image[ch * N + i] = g;
}

Win11 make_window()
{
	// utils/loss_utils.py:359-361: exp(-(x - 5)^2 / (2 sigma^2)) as fp32 values, normalised in fp32
	Win11 w;
	float s = 0.f;
	for (int i = 0; i < 11; i++) { w.w[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); s += w.w[i]; }
	for (int i = 0; i < 11; i++) w.w[i] /= s;
	return w;
}

template <bool GRAY, typename GT>
void launch_both(int W, int H, const float* img, const void* gt, float lambda, float upstream, float* maps, double* sums,
	float* d_image, float* out3, cudaStream_t st)
{
	const Win11 win = make_window();
	const dim3 grid((W + SSIM_T - 1) / SSIM_T, (H + SSIM_T - 1) / SSIM_T, GRAY ? 1 : 3), block(SSIM_T, SSIM_T);
	ssim_fwd_kernel<GRAY, GT><<<grid, block, 0, st>>>(W, H, img, (const GT*)gt, win, maps, sums);
	ssim_bwd_kernel<GRAY, GT><<<grid, block, 0, st>>>(W, H, img, (const GT*)gt, win, maps, sums, lambda, upstream, d_image, out3);
}

}  // namespace

size_t e3_l1_ssim_scratch_words(int W, int H, bool gray) { return 16 + (size_t)3 * (gray ? 1 : 3) * W * H; }

cudaError_t e3_launch_l1_ssim(int W, int H, bool gray, const float* img, const void* gt, bool gt_is_u8, float lambda,
	float upstream, float* scratch, float* d_image, float* out3, cudaStream_t st)
{
	double* sums = reinterpret_cast<double*>(scratch);
	float* maps = scratch + 16;
	cudaError_t e = cudaMemsetAsync(sums, 0, 2 * sizeof(double), st);
	if (e != cudaSuccess) return e;
	if (gray) { if (gt_is_u8) launch_both<true, uint8_t>(W, H, img, gt, lambda, upstream, maps, sums, d_image, out3, st);
		else launch_both<true, float>(W, H, img, gt, lambda, upstream, maps, sums, d_image, out3, st); }
	else { if (gt_is_u8) launch_both<false, uint8_t>(W, H, img, gt, lambda, upstream, maps, sums, d_image, out3, st);
		else launch_both<false, float>(W, H, img, gt, lambda, upstream, maps, sums, d_image, out3, st); }
	g_e3_launches += 2;
	return cudaGetLastError();
}
