// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// C-ABI shim around the *unmodified* reference rasterizer
// (CudaRasterizer::Rasterizer, /root/reference/submodules/diff-gaussian-rasterization/
// cuda_rasterizer/rasterizer.h:20-85). The reference sources are compiled where they lie
// by oracle/Makefile into oracle/_ref/libref_dgr.so; this file replaces only the torch glue
// (rasterize_points.cu:35-196) with a buffer pool so that it can be driven through ctypes.
// Used by tests/ (bit-exact parity target on the GPU box) and by bench.py's reference-CUDA
// timing leg. It needs a GPU to run; /root/reference is needed only at build time.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <cuda_runtime.h>
#include "cuda_rasterizer/config.h"
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {
struct Pool {
	char* ptr = nullptr;
	size_t cap = 0;
	size_t used = 0;
	char* get(size_t n) {
		if (n > cap) {
			if (ptr) cudaFree(ptr);
			size_t want = n + n / 4 + 256;
			if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
			cap = want;
		}
		used = n;
		return ptr;
	}
	void release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = used = 0; }
};
struct RefHandle {
	Pool geom, binning, img;
	int P = 0, R = 0, W = 0, H = 0;
};
}

extern "C" {

void* ref_create() { return new RefHandle(); }

void ref_destroy(void* h) {
	RefHandle* r = (RefHandle*)h;
	if (!r) return;
	r->geom.release(); r->binning.release(); r->img.release();
	delete r;
}

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-115). Null pointers stand for the
// empty tensors of the python side. Returns num_rendered, or -1 on a CUDA error.
int ref_forward(void* h, int P, int D, int M,
	const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* opacities, const float* scales, float scale_modifier,
	const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, int* radii, int debug)
{
	RefHandle* r = (RefHandle*)h;
	r->P = P; r->W = W; r->H = H; r->R = 0;
	if (P == 0) return 0;
	auto g = [r](size_t n) { return r->geom.get(n); };
	auto b = [r](size_t n) { return r->binning.get(n); };
	auto i = [r](size_t n) { return r->img.get(n); };
	try {
		r->R = CudaRasterizer::Rasterizer::forward(g, b, i, P, D, M, background, W, H,
			means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations,
			cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
			prefiltered != 0, out_color, radii, debug != 0);
	} catch (const std::exception& e) {
		fprintf(stderr, "ref_forward: %s\n", e.what());
		return -1;
	}
	return r->R;
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:117-196). The nine gradient
// arrays must be zero-filled by the caller (the reference does that with torch::zeros).
int ref_backward(void* h, int P, int D, int M, int R,
	const float* background, int W, int H,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations,
	const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
	const float* campos, float tan_fovx, float tan_fovy, const int* radii,
	const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
	float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
	float* dL_dscale, float* dL_drot, int debug)
{
	RefHandle* r = (RefHandle*)h;
	if (P == 0) return 0;
	try {
		CudaRasterizer::Rasterizer::backward(P, D, M, R, background, W, H, means3D, shs,
			colors_precomp, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
			projmatrix, campos, tan_fovx, tan_fovy, radii, r->geom.ptr, r->binning.ptr,
			r->img.ptr, dL_dpix, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D,
			dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
	} catch (const std::exception& e) {
		fprintf(stderr, "ref_backward: %s\n", e.what());
		return -1;
	}
	return 0;
}

int ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present)
{
	if (P) CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
	return 0;
}

// Device pointers into the opaque state of the last forward (layout: rasterizer_impl.cu:155-194).
// which: 0 depths f32[P], 1 means2D f32[2P], 2 cov3D f32[6P], 3 conic_opacity f32[4P],
// 4 tiles_touched u32[P], 5 point_offsets u32[P], 6 point_list u32[R], 7 point_list_keys u64[R],
// 8 ranges u32[2T], 9 n_contrib u32[N], 10 accum_alpha f32[N], 11 keys_unsorted u64[R],
// 12 point_list_unsorted u32[R], 13 rgb f32[3P], 14 clamped u8[3P]
void* ref_state_ptr(void* h, int which)
{
	RefHandle* r = (RefHandle*)h;
	if (!r->geom.ptr) return nullptr;
	char* c = r->geom.ptr;
	CudaRasterizer::GeometryState geo = CudaRasterizer::GeometryState::fromChunk(c, r->P);
	c = r->img.ptr;
	CudaRasterizer::ImageState img = CudaRasterizer::ImageState::fromChunk(c, (size_t)r->W * r->H);
	CudaRasterizer::BinningState bin{};
	if (r->binning.ptr) { c = r->binning.ptr; bin = CudaRasterizer::BinningState::fromChunk(c, r->R); }
	switch (which) {
	case 0: return geo.depths;
	case 1: return geo.means2D;
	case 2: return geo.cov3D;
	case 3: return geo.conic_opacity;
	case 4: return geo.tiles_touched;
	case 5: return geo.point_offsets;
	case 6: return bin.point_list;
	case 7: return bin.point_list_keys;
	case 8: return img.ranges;
	case 9: return img.n_contrib;
	case 10: return img.accum_alpha;
	case 11: return bin.point_list_keys_unsorted;
	case 12: return bin.point_list_unsorted;
	case 13: return geo.rgb;
	case 14: return geo.clamped;
	}
	return nullptr;
}

// Copy `nbytes` of state array `which` into a caller-owned device buffer.
int ref_copy_state(void* h, int which, void* dst, size_t nbytes)
{
	void* src = ref_state_ptr(h, which);
	if (!src) return -1;
	cudaDeviceSynchronize();
	return (int)cudaMemcpy(dst, src, nbytes, cudaMemcpyDeviceToDevice);
}

}
