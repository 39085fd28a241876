// Event loss: forward value and backward seeds in one streaming pass over the rendered images. sm_100a.
//
// Reference: train.py:165-203 with utils/loss_utils.py:24-28 (rgb_to_LUVscale), :234-249
// (differentialable_event_simu) and :270-271 (l1_loss). There it is ~50 PyTorch launches over
// N-sized tensors plus their autograd backward, and an ssim_gray evaluation that is multiplied by
// lambda_dssim = 0 (train.py:177-178). Here:
//   pre-pass (event_gt_fraction_kernel): f = the fraction of pixels whose GT log-luma difference is not zero. It
//           weighs the two loss terms (train.py:190-196) and depends on the two GT frames alone, so it can run
//           while the views are still being rendered;
//   main pass (event_loss_kernel): per pixel the predicted and the GT log-luma difference, the closed-form seeds
//           dL/dimage, dL/dimage_now, dL/dimage_next (their factors need f only), and four
//           sums (fp64 atomics: |d-g|, |image-gt|, |image-blur|, sgn(d-g)*d) from which the block that finishes
//           last writes the loss value and dL/dc.
// Four pixels per thread (16-byte loads and stores, 4-byte loads of 8-bit GT) when N % 4 == 0.
// Algorithmic bytes of the main pass: 6 (7 with --deblur) [3,H,W] images in, 3 seeds out = 108 (120) B/pixel with
// fp32 GT frames and RGB event views; the pre-pass reads 24 B/pixel. HBM-bound streaming kernels.
#include "common.cuh"
#include <initializer_list>

namespace {

#define LUMA_R 0.4124f
#define LUMA_G 0.35758f
#define LUMA_B 0.1804f
#define EVT_EPS 1e-8f

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

// ---- V pixels starting at i: loads of the image planes ----
template <int V> struct Px { float v[V]; };
template <int V> __device__ __forceinline__ Px<V> ld_plane(const float* __restrict__ p, size_t i)
{
	Px<V> r;
	if constexpr (V == 4) { const float4 t = __ldg(reinterpret_cast<const float4*>(p + i)); r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w; }
	else r.v[0] = __ldg(p + i);
	return r;
}
template <int V> __device__ __forceinline__ Px<V> ld_plane(const uint8_t* __restrict__ p, size_t i)
{
	Px<V> r;
	// GT frames may stay in the 8-bit form they have on disk (the reference converts them to fp32 / 255 once at load
	// time, utils/general_utils.py:21-27, and keeps 12 bytes per pixel resident)
	if constexpr (V == 4) {
		const uchar4 t = __ldg(reinterpret_cast<const uchar4*>(p + i));
		r.v[0] = (float)t.x / 255.0f; r.v[1] = (float)t.y / 255.0f; r.v[2] = (float)t.z / 255.0f; r.v[3] = (float)t.w / 255.0f;
	} else r.v[0] = (float)__ldg(p + i) / 255.0f;
	return r;
}
template <int V> __device__ __forceinline__ void st_plane(float* __restrict__ p, size_t i, const Px<V>& a)
{
	if constexpr (V == 4) *reinterpret_cast<float4*>(p + i) = make_float4(a.v[0], a.v[1], a.v[2], a.v[3]);
	else p[i] = a.v[0];
}
// luma of V pixels of a [C,N] image (C = 1: the plane itself); EC = channels of the two event renders: 3 (RGB, luma
// taken here) or 1 (already luma)
template <int V, int C, typename T> __device__ __forceinline__ Px<V> luma_px(const T* __restrict__ img, size_t i, size_t N)
{
	if (C == 1) return ld_plane<V>(img, i);
	const Px<V> r = ld_plane<V>(img, i), g = ld_plane<V>(img, N + i), b = ld_plane<V>(img, 2 * N + i);
	Px<V> y;
#pragma unroll
	for (int k = 0; k < V; k++) y.v[k] = LUMA_R * r.v[k] + LUMA_G * g.v[k] + LUMA_B * b.v[k];
	return y;
}
__device__ __forceinline__ float gt_event(float y_now, float y_next) { return (logf(y_next + EVT_EPS) - logf(y_now + EVT_EPS)) / 0.17f; }

// sums[1] = #(g != 0), g = the GT log-luma difference of the two event frames
template <int V, typename GT>
__global__ void __launch_bounds__(256) event_gt_fraction_kernel(size_t N, const GT* __restrict__ gt_now, const GT* __restrict__ gt_next,
	double* __restrict__ sums)
{
	__shared__ double sm[8];
	double s1 = 0;
	for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * V; i < N; i += (size_t)gridDim.x * blockDim.x * V) {
		const Px<V> a = luma_px<V, 3>(gt_now, i, N), b = luma_px<V, 3>(gt_next, i, N);
#pragma unroll
		for (int k = 0; k < V; k++) s1 += (gt_event(a.v[k], b.v[k]) != 0.f) ? 1.0 : 0.0;
	}
	s1 = block_sum(s1, sm);
	if (threadIdx.x == 0) atomicAdd(sums + 1, s1);
}

// sums[0] = sum |d - g|, [1] = #(g != 0) (from the pre-pass), [2] = sum |image - gt_int| (3N terms),
// [3] = sum |image - gt_blur|, [4] = sum sgn(d - g) * d; the word at sums + 6 counts the finished blocks.
// out[0] = loss, out[1] = dL/dc, out[2] = loss1 (event L1), out[3] = loss2 (intensity L1),
// out[4] = fraction of pixels with a non-zero GT event difference
// TOT (shared first view: image_now IS image, rendered once): d_now receives the TOTAL seed of that view,
// event-now term + intensity term, while d_image keeps the intensity term alone (the second seed of the
// dual backward blend)
template <int V, int ECN, int ECX, bool TOT, typename GT>
__global__ void __launch_bounds__(256) event_loss_kernel(size_t N, const float* __restrict__ image,
	const float* __restrict__ image_now, const float* __restrict__ image_next, const GT* __restrict__ gt_now,
	const GT* __restrict__ gt_next, const GT* __restrict__ gt_int, const GT* __restrict__ gt_blur,
	const float* __restrict__ c_ptr, double* __restrict__ sums, float upstream, float* __restrict__ d_image,
	float* __restrict__ d_now, float* __restrict__ d_next, float* __restrict__ out)
{
	__shared__ double sm[8];
	__shared__ bool s_last;
	const float c = __ldg(c_ptr);
	const double Nd = (double)N;
	const double f = sums[1] / Nd;
	const double scale = gt_blur ? 0.5 : 1.0;                       // train.py:201-202
	const float k1 = (float)(upstream * scale * 0.9 * f / Nd);      // event term, per pixel
	const float k2 = (float)(upstream * scale * 0.1 * (1.0 - f) / (3.0 * Nd));   // intensity term, per element
	const float k3 = gt_blur ? (float)(upstream * 0.5 / (3.0 * Nd)) : 0.f;       // blurry-frame term
	double s0 = 0, s2 = 0, s3 = 0, s4 = 0;
	for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * V; i < N; i += (size_t)gridDim.x * blockDim.x * V) {
		const Px<V> y1 = luma_px<V, ECN>(image_now, i, N), y2 = luma_px<V, ECX>(image_next, i, N);
		const Px<V> g1 = luma_px<V, 3>(gt_now, i, N), g2 = luma_px<V, 3>(gt_next, i, N);
		Px<V> dy1, dy2;
#pragma unroll
		for (int k = 0; k < V; k++) {
			const float d = (logf(y2.v[k] + EVT_EPS) - logf(y1.v[k] + EVT_EPS)) / c;
			const float g = gt_event(g1.v[k], g2.v[k]);
			const float s = sgn(d - g);
			s0 += fabsf(d - g);
			s4 += s * d;
			dy2.v[k] = k1 * s / (c * (y2.v[k] + EVT_EPS));
			dy1.v[k] = -k1 * s / (c * (y1.v[k] + EVT_EPS));
		}
		if (ECX == 1) st_plane<V>(d_next, i, dy2);
		else {
			Px<V> t;
#pragma unroll
			for (int k = 0; k < V; k++) t.v[k] = LUMA_R * dy2.v[k];
			st_plane<V>(d_next, i, t);
#pragma unroll
			for (int k = 0; k < V; k++) t.v[k] = LUMA_G * dy2.v[k];
			st_plane<V>(d_next, N + i, t);
#pragma unroll
			for (int k = 0; k < V; k++) t.v[k] = LUMA_B * dy2.v[k];
			st_plane<V>(d_next, 2 * N + i, t);
		}
		Px<V> gi[3];
#pragma unroll
		for (int ch = 0; ch < 3; ch++) {
			const Px<V> v = ld_plane<V>(image, ch * N + i), t = ld_plane<V>(gt_int, ch * N + i);
			Px<V> gr;
#pragma unroll
			for (int k = 0; k < V; k++) { s2 += fabsf(v.v[k] - t.v[k]); gr.v[k] = k2 * sgn(v.v[k] - t.v[k]); }
			if (gt_blur) {
				const Px<V> bl = ld_plane<V>(gt_blur, ch * N + i);
#pragma unroll
				for (int k = 0; k < V; k++) { s3 += fabsf(v.v[k] - bl.v[k]); gr.v[k] += k3 * sgn(v.v[k] - bl.v[k]); }
			}
			st_plane<V>(d_image, ch * N + i, gr);
#pragma unroll
			for (int k = 0; k < V; k++) gi[ch].v[k] = TOT ? gr.v[k] : 0.f;
		}
		if (ECN == 1) st_plane<V>(d_now, i, dy1);
		else {
			const float w[3] = { LUMA_R, LUMA_G, LUMA_B };
#pragma unroll
			for (int ch = 0; ch < 3; ch++) {
				Px<V> t;
#pragma unroll
				for (int k = 0; k < V; k++) t.v[k] = w[ch] * dy1.v[k] + gi[ch].v[k];
				st_plane<V>(d_now, ch * N + i, t);
			}
		}
	}
	s0 = block_sum(s0, sm); s2 = block_sum(s2, sm); s3 = block_sum(s3, sm); s4 = block_sum(s4, sm);
	if (threadIdx.x == 0) {
		atomicAdd(sums + 0, s0); atomicAdd(sums + 2, s2); atomicAdd(sums + 3, s3); atomicAdd(sums + 4, s4);
		__threadfence();
		s_last = atomicAdd(reinterpret_cast<unsigned int*>(sums + 6), 1u) == gridDim.x - 1;
	}
	__syncthreads();
	if (s_last && threadIdx.x == 0) {
		__threadfence();
		volatile double* vs = sums;
		const double loss1 = vs[0] / Nd, loss2 = vs[2] / (3.0 * Nd);
		double loss = 0.9 * loss1 * f + 0.1 * loss2 * (1.0 - f);     // train.py:190-196
		if (gt_blur) loss = 0.5 * loss + 0.5 * (vs[3] / (3.0 * Nd));
		out[0] = (float)loss;
		out[1] = (float)(-(double)upstream * scale * 0.9 * f / Nd * vs[4] / (double)c);
		out[2] = (float)loss1; out[3] = (float)loss2; out[4] = (float)f;
	}
}

}  // namespace

static bool vec4_ok(size_t N, std::initializer_list<const void*> ptrs, bool gt_is_u8, std::initializer_list<const void*> gts)
{
	if (N % 4) return false;
	for (const void* p : ptrs) if (reinterpret_cast<size_t>(p) & 15) return false;
	for (const void* p : gts) if (p && (reinterpret_cast<size_t>(p) & (gt_is_u8 ? 3 : 15))) return false;
	return true;
}

// sums <- 0, then sums[1] = #(pixels with a non-zero GT event difference); the first half of the event loss, needs the GT alone
cudaError_t e3_launch_event_gt_fraction(size_t N, const void* gt_now, const void* gt_next, int gt_is_u8, double* sums, cudaStream_t st)
{
	cudaError_t e = cudaMemsetAsync(sums, 0, 8 * sizeof(double), st);
	if (e != cudaSuccess) return e;
	const bool v4 = vec4_ok(N, {}, gt_is_u8 != 0, { gt_now, gt_next });
	const size_t per = v4 ? 4 : 1;
	const int grid = (int)min((size_t)148 * 8, (N / per + 255) / 256 + 1);
	if (gt_is_u8) {
		if (v4) event_gt_fraction_kernel<4, uint8_t><<<grid, 256, 0, st>>>(N, (const uint8_t*)gt_now, (const uint8_t*)gt_next, sums);
		else event_gt_fraction_kernel<1, uint8_t><<<grid, 256, 0, st>>>(N, (const uint8_t*)gt_now, (const uint8_t*)gt_next, sums);
	} else {
		if (v4) event_gt_fraction_kernel<4, float><<<grid, 256, 0, st>>>(N, (const float*)gt_now, (const float*)gt_next, sums);
		else event_gt_fraction_kernel<1, float><<<grid, 256, 0, st>>>(N, (const float*)gt_now, (const float*)gt_next, sums);
	}
	g_e3_launches += 1;
	return cudaGetLastError();
}

template <int ECN, int ECX, bool TOT, typename GT>
static void launch_loss(bool v4, size_t N, const float* image, const float* image_now, const float* image_next, const void* gt_now,
	const void* gt_next, const void* gt_int, const void* gt_blur, const float* c_ptr, float upstream, double* sums, float* d_image,
	float* d_now, float* d_next, float* out, cudaStream_t st)
{
	const size_t per = v4 ? 4 : 1;
	const int grid = (int)min((size_t)148 * 8, (N / per + 255) / 256 + 1);
	if (v4)
		event_loss_kernel<4, ECN, ECX, TOT, GT><<<grid, 256, 0, st>>>(N, image, image_now, image_next, (const GT*)gt_now, (const GT*)gt_next,
			(const GT*)gt_int, (const GT*)gt_blur, c_ptr, sums, upstream, d_image, d_now, d_next, out);
	else
		event_loss_kernel<1, ECN, ECX, TOT, GT><<<grid, 256, 0, st>>>(N, image, image_now, image_next, (const GT*)gt_now, (const GT*)gt_next,
			(const GT*)gt_int, (const GT*)gt_blur, c_ptr, sums, upstream, d_image, d_now, d_next, out);
}

// The second half: needs e3_launch_event_gt_fraction on the same `sums` before it (stream-ordered).
// event_channels: 1 (both event renders are luma), 3 (both RGB), or E3_EVENT_SHARED_FIRST_VIEW (image_now is the
// 3-channel intensity render itself, image_next a luma render; d_now = total seed of the shared view)
cudaError_t e3_launch_event_loss(size_t N, const float* image, const float* image_now, const float* image_next,
	const void* gt_now, const void* gt_next, const void* gt_int, const void* gt_blur, int gt_is_u8, const float* c_ptr,
	float upstream, int event_channels, double* sums, float* d_image, float* d_now, float* d_next, float* out, cudaStream_t st)
{
	const bool v4 = vec4_ok(N, { image, image_now, image_next, d_image, d_now, d_next }, gt_is_u8 != 0, { gt_now, gt_next, gt_int, gt_blur });
#define ARGS v4, N, image, image_now, image_next, gt_now, gt_next, gt_int, gt_blur, c_ptr, upstream, sums, d_image, d_now, d_next, out, st
	if (event_channels == 1) { if (gt_is_u8) launch_loss<1, 1, false, uint8_t>(ARGS); else launch_loss<1, 1, false, float>(ARGS); }
	else if (event_channels == 3) { if (gt_is_u8) launch_loss<3, 3, false, uint8_t>(ARGS); else launch_loss<3, 3, false, float>(ARGS); }
	else { if (gt_is_u8) launch_loss<3, 1, true, uint8_t>(ARGS); else launch_loss<3, 1, true, float>(ARGS); }
#undef ARGS
	g_e3_launches += 1;
	return cudaGetLastError();
}
