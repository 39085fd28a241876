// Event loss, forward value and backward seeds in two streaming passes. sm_100a.
//
// Reference: train.py:165-203 with utils/loss_utils.py:24-28 (rgb_to_LUVscale), :234-249
// (differentialable_event_simu) and :270-271 (l1_loss). There it is ~50 PyTorch launches over
// N-sized tensors plus their autograd backward, and an ssim_gray evaluation that is multiplied by
// lambda_dssim = 0 (train.py:177-178). Here:
//   pass 1 (event_loss_reduce_kernel): per pixel the predicted and the GT log-luma difference,
//           accumulated into 5 sums (fp64 atomics: |d-g|, #(g != 0), |image-gt|, |image-blur|,
//           sgn(d-g)*d);
//   pass 2 (event_loss_seed_kernel): the closed-form seeds dL/dimage, dL/dimage_now,
//           dL/dimage_next, the loss value and dL/dc.
// Algorithmic bytes: pass 1 reads 6 (7 with --deblur) [3,H,W] images = 72 (84) B/pixel, pass 2
// reads them again and writes 3 seeds = 108 (120) B/pixel more. HBM-bound streaming kernels.
#include "common.cuh"

namespace {

#define LUMA_R 0.4124f
#define LUMA_G 0.35758f
#define LUMA_B 0.1804f
#define EVT_EPS 1e-8f

// EC = channels of the two event renders: 3 (RGB, luma taken here) or 1 (already luma)
template <int EC>
__device__ __forceinline__ float luma_ev(const float* __restrict__ img, size_t i, size_t N)
{
	if (EC == 1) return __ldg(img + i);
	return LUMA_R * __ldg(img + i) + LUMA_G * __ldg(img + N + i) + LUMA_B * __ldg(img + 2 * N + i);
}

// GT frames may stay in the 8-bit form they have on disk (the reference converts them to fp32 /255
// once at load time, utils/general_utils.py:21-27, and keeps 12 bytes per pixel resident)
template <typename GT> __device__ __forceinline__ float gt_load(const GT* __restrict__ p, size_t i);
template <> __device__ __forceinline__ float gt_load<float>(const float* __restrict__ p, size_t i) { return __ldg(p + i); }
template <> __device__ __forceinline__ float gt_load<uint8_t>(const uint8_t* __restrict__ p, size_t i) { return (float)__ldg(p + i) / 255.0f; }

template <typename GT>
__device__ __forceinline__ float luma_gt(const GT* __restrict__ img, size_t i, size_t N)
{
	return LUMA_R * gt_load<GT>(img, i) + LUMA_G * gt_load<GT>(img, N + i) + LUMA_B * gt_load<GT>(img, 2 * N + i);
}


__device__ __forceinline__ float sgn(float v) { return (v > 0.f) - (v < 0.f); }

__device__ __forceinline__ double block_sum(double v, double* sm)
{
	for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(E3_FULL, v, o);
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	__syncthreads();
	if (lane == 0) sm[w] = v;
	__syncthreads();
	double r = 0.0;
	if (threadIdx.x == 0)
		for (int i = 0; i < (int)(blockDim.x >> 5); i++) r += sm[i];
	return r;
}

// sums[0] = sum |d - g|, [1] = #(g != 0), [2] = sum |image - gt_int| (3N terms),
// [3] = sum |image - gt_blur|, [4] = sum sgn(d - g) * d
template <int ECN, int ECX, typename GT>
__global__ void __launch_bounds__(256) event_loss_reduce_kernel(size_t N, const float* __restrict__ image,
	const float* __restrict__ image_now, const float* __restrict__ image_next, const GT* __restrict__ gt_now,
	const GT* __restrict__ gt_next, const GT* __restrict__ gt_int, const GT* __restrict__ gt_blur,
	const float* __restrict__ c_ptr, double* __restrict__ sums)
{
	__shared__ double sm[8];
	const float c = __ldg(c_ptr);
	double s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (size_t)gridDim.x * blockDim.x) {
		const float y1 = luma_ev<ECN>(image_now, i, N), y2 = luma_ev<ECX>(image_next, i, N);
		const float d = (logf(y2 + EVT_EPS) - logf(y1 + EVT_EPS)) / c;
		const float g = (logf(luma_gt<GT>(gt_next, i, N) + EVT_EPS) - logf(luma_gt<GT>(gt_now, i, N) + EVT_EPS)) / 0.17f;
		s0 += fabsf(d - g);
		s1 += (g != 0.f) ? 1.0 : 0.0;
		s4 += sgn(d - g) * d;
#pragma unroll
		for (int ch = 0; ch < 3; ch++) {
			const float v = __ldg(image + ch * N + i);
			s2 += fabsf(v - gt_load<GT>(gt_int, ch * N + i));
			if (gt_blur) s3 += fabsf(v - gt_load<GT>(gt_blur, ch * N + i));
		}
	}
	s0 = block_sum(s0, sm); s1 = block_sum(s1, sm); s2 = block_sum(s2, sm);
	s3 = block_sum(s3, sm); s4 = block_sum(s4, sm);
	if (threadIdx.x == 0) {
		atomicAdd(sums + 0, s0); atomicAdd(sums + 1, s1); atomicAdd(sums + 2, s2);
		atomicAdd(sums + 3, s3); atomicAdd(sums + 4, s4);
	}
}

// out[0] = loss, out[1] = dL/dc, out[2] = loss1 (event L1), out[3] = loss2 (intensity L1),
// out[4] = fraction of pixels with a non-zero GT event difference
// TOT (shared first view: image_now IS image, rendered once): d_now receives the TOTAL seed of that view,
// event-now term + intensity term, while d_image keeps the intensity term alone (the second seed of the
// dual backward blend)
template <int ECN, int ECX, bool TOT, typename GT>
__global__ void __launch_bounds__(256) event_loss_seed_kernel(size_t N, const float* __restrict__ image,
	const float* __restrict__ image_now, const float* __restrict__ image_next, const GT* __restrict__ gt_now,
	const GT* __restrict__ gt_next, const GT* __restrict__ gt_int, const GT* __restrict__ gt_blur,
	const float* __restrict__ c_ptr, const double* __restrict__ sums, float upstream, float* __restrict__ d_image,
	float* __restrict__ d_now, float* __restrict__ d_next, float* __restrict__ out)
{
	const float c = __ldg(c_ptr);
	const double Nd = (double)N;
	const double f = sums[1] / Nd;
	const double scale = gt_blur ? 0.5 : 1.0;                       // train.py:201-202
	const float k1 = (float)(upstream * scale * 0.9 * f / Nd);      // event term, per pixel
	const float k2 = (float)(upstream * scale * 0.1 * (1.0 - f) / (3.0 * Nd));   // intensity term, per element
	const float k3 = gt_blur ? (float)(upstream * 0.5 / (3.0 * Nd)) : 0.f;       // blurry-frame term
	if (blockIdx.x == 0 && threadIdx.x == 0) {
		const double loss1 = sums[0] / Nd, loss2 = sums[2] / (3.0 * Nd);
		double loss = 0.9 * loss1 * f + 0.1 * loss2 * (1.0 - f);     // train.py:190-196
		if (gt_blur) loss = 0.5 * loss + 0.5 * (sums[3] / (3.0 * Nd));
		out[0] = (float)loss;
		out[1] = (float)(-(double)upstream * scale * 0.9 * f / Nd * sums[4] / (double)c);
		out[2] = (float)loss1; out[3] = (float)loss2; out[4] = (float)f;
	}
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += (size_t)gridDim.x * blockDim.x) {
		const float y1 = luma_ev<ECN>(image_now, i, N), y2 = luma_ev<ECX>(image_next, i, N);
		const float d = (logf(y2 + EVT_EPS) - logf(y1 + EVT_EPS)) / c;
		const float g = (logf(luma_gt<GT>(gt_next, i, N) + EVT_EPS) - logf(luma_gt<GT>(gt_now, i, N) + EVT_EPS)) / 0.17f;
		const float s = sgn(d - g);
		const float dy2 = k1 * s / (c * (y2 + EVT_EPS));
		const float dy1 = -k1 * s / (c * (y1 + EVT_EPS));
		if (ECX == 1) d_next[i] = dy2;
		else { d_next[i] = LUMA_R * dy2; d_next[N + i] = LUMA_G * dy2; d_next[2 * N + i] = LUMA_B * dy2; }
		float gi[3];
#pragma unroll
		for (int ch = 0; ch < 3; ch++) {
			const float v = __ldg(image + ch * N + i);
			float gr = k2 * sgn(v - gt_load<GT>(gt_int, ch * N + i));
			if (gt_blur) gr += k3 * sgn(v - gt_load<GT>(gt_blur, ch * N + i));
			d_image[ch * N + i] = gr;
			gi[ch] = TOT ? gr : 0.f;
		}
		if (ECN == 1) d_now[i] = dy1;
		else { d_now[i] = LUMA_R * dy1 + gi[0]; d_now[N + i] = LUMA_G * dy1 + gi[1]; d_now[2 * N + i] = LUMA_B * dy1 + gi[2]; }
	}
}

}  // namespace

template <int ECN, int ECX, bool TOT, typename GT>
static void launch_loss(int grid, size_t N, const float* image, const float* image_now, const float* image_next, const void* gt_now,
	const void* gt_next, const void* gt_int, const void* gt_blur, const float* c_ptr, float upstream, double* sums, float* d_image,
	float* d_now, float* d_next, float* out, cudaStream_t st)
{
	event_loss_reduce_kernel<ECN, ECX, GT><<<grid, 256, 0, st>>>(N, image, image_now, image_next, (const GT*)gt_now, (const GT*)gt_next,
		(const GT*)gt_int, (const GT*)gt_blur, c_ptr, sums);
	event_loss_seed_kernel<ECN, ECX, TOT, GT><<<grid, 256, 0, st>>>(N, image, image_now, image_next, (const GT*)gt_now, (const GT*)gt_next,
		(const GT*)gt_int, (const GT*)gt_blur, c_ptr, sums, upstream, d_image, d_now, d_next, out);
}

// event_channels: 1 (both event renders are luma), 3 (both RGB), or E3_EVENT_SHARED_FIRST_VIEW (image_now is the
// 3-channel intensity render itself, image_next a luma render; d_now = total seed of the shared view)
cudaError_t e3_launch_event_loss(size_t N, const float* image, const float* image_now, const float* image_next,
	const void* gt_now, const void* gt_next, const void* gt_int, const void* gt_blur, int gt_is_u8, const float* c_ptr,
	float upstream, int event_channels, double* sums, float* d_image, float* d_now, float* d_next, float* out, cudaStream_t st)
{
	cudaError_t e = cudaMemsetAsync(sums, 0, 8 * sizeof(double), st);
	if (e != cudaSuccess) return e;
	const int grid = (int)min((size_t)148 * 8, (N + 255) / 256);
#define ARGS grid, N, image, image_now, image_next, gt_now, gt_next, gt_int, gt_blur, c_ptr, upstream, sums, d_image, d_now, d_next, out, st
	if (event_channels == 1) { if (gt_is_u8) launch_loss<1, 1, false, uint8_t>(ARGS); else launch_loss<1, 1, false, float>(ARGS); }
	else if (event_channels == 3) { if (gt_is_u8) launch_loss<3, 3, false, uint8_t>(ARGS); else launch_loss<3, 3, false, float>(ARGS); }
	else { if (gt_is_u8) launch_loss<3, 1, true, uint8_t>(ARGS); else launch_loss<3, 1, true, float>(ARGS); }
#undef ARGS
	g_e3_launches += 2;
	return cudaGetLastError();
}
