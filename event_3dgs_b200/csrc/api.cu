// extern "C" entry points of libevent3dgs.so (declared in include/event3dgs.h).
// Host orchestration mirrors CudaRasterizer::Rasterizer::forward / backward
// (cuda_rasterizer/rasterizer_impl.cu:198-336, 340-434).
#include "common.cuh"
#include "../../include/event3dgs.h"

#include <cstdio>
#include <mutex>
#include <set>
#include <utility>
#include <vector>

std::atomic<unsigned long long> g_e3_launches{0};

cudaError_t e3_allow_dynamic_smem_impl(const void* kernel, size_t bytes)
{
	static std::mutex mu;
	static std::set<std::pair<const void*, int>> done;
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return e;
	std::lock_guard<std::mutex> l(mu);
	if (done.count({ kernel, dev })) return cudaSuccess;
	e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
	if (e == cudaSuccess) done.insert({ kernel, dev });
	return e;
}

namespace {

// rasterizer_impl.cu:35-50: number of tile-id bits the sort has to cover
uint32_t higher_msb(uint32_t n)
{
	uint32_t msb = sizeof(n) * 4, step = msb;
	while (step > 1) {
		step /= 2;
		if (n >> msb) msb += step; else msb -= step;
	}
	if (n >> msb) msb++;
	return msb;
}

E3Frame make_frame(int P, int D, int M, int C, int W, int H, float tan_fovx, float tan_fovy, float mod)
{
	E3Frame f;
	f.P = P; f.D = D; f.M = M; f.C = C; f.W = W; f.H = H;
	f.gx = (W + E3_TILE_X - 1) / E3_TILE_X;
	f.gy = (H + E3_TILE_Y - 1) / E3_TILE_Y;
	f.tan_fovx = tan_fovx; f.tan_fovy = tan_fovy;
	f.focal_y = H / (2.0f * tan_fovy);     // rasterizer_impl.cu:222-223
	f.focal_x = W / (2.0f * tan_fovx);
	f.scale_modifier = mod;
	return f;
}

#define E3_CHECK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return (int)e__; } while (0)

// ---- optional per-stage device timing (CUDA events on the launching stream) ----
struct ProfSpan { int stage; cudaEvent_t a, b; };
unsigned g_prof_mask = 0;
std::vector<ProfSpan> g_prof_open;
std::vector<cudaEvent_t> g_prof_pool;
double g_prof_ms[E3_NUM_STAGES];
long long g_prof_n[E3_NUM_STAGES];
std::mutex g_prof_mu;

cudaEvent_t prof_event()
{
	cudaEvent_t e;
	if (!g_prof_pool.empty()) { e = g_prof_pool.back(); g_prof_pool.pop_back(); return e; }
	cudaEventCreate(&e);
	return e;
}

struct ProfScope {
	int stage; cudaStream_t st; bool on; cudaEvent_t a;
	ProfScope(int stage_, cudaStream_t st_) : stage(stage_), st(st_), on((g_prof_mask >> stage_) & 1u)
	{
		if (on) { std::lock_guard<std::mutex> l(g_prof_mu); a = prof_event(); cudaEventRecord(a, st); }
	}
	~ProfScope()
	{
		if (on) { std::lock_guard<std::mutex> l(g_prof_mu); cudaEvent_t b = prof_event(); cudaEventRecord(b, st); g_prof_open.push_back({ stage, a, b }); }
	}
};

// debug mode: sync-and-check after every stage, like CHECK_CUDA (auxiliary.h:166-173)
#define E3_STAGE(id, expr) do { cudaError_t es__; { ProfScope ps__(id, st); es__ = (expr); } \
	if (es__ != cudaSuccess) return (int)es__; if (debug) { E3_CHECK(cudaStreamSynchronize(st)); } } while (0)

}  // namespace

extern "C" {

int e3_version(void) { return 100; }

const char* e3_error_string(int code)
{
	switch (code) {
	case 0: return "success";
	case E3_ERR_BAD_ARGUMENT: return "event3dgs: bad argument";
	case E3_ERR_ALLOC_FAILED: return "event3dgs: workspace allocation callback returned NULL";
	case E3_ERR_PREFILTERED: return "Point is filtered although prefiltered is set. This shouldn't happen!";
	case E3_ERR_UNSUPPORTED: return "For non-RGB, provide precomputed Gaussian colors!";
	case E3_ERR_TOO_LARGE: return "event3dgs: input larger than this entry point's fixed capacity (e3_exclusive_scan_noalloc: 16777216 values; e3_densify_plan: 2^30 rows)";
	}
	if (code > 0) return cudaGetErrorString((cudaError_t)code);
	return "event3dgs: unknown error";
}

size_t e3_geometry_bytes(int P)
{
	char* end = nullptr;
	E3GeomState::carve(nullptr, (size_t)P, &end);
	return (size_t)end + 256;
}

size_t e3_image_bytes(int width, int height)
{
	char* end = nullptr;
	const size_t T = (size_t)((width + E3_TILE_X - 1) / E3_TILE_X) * ((height + E3_TILE_Y - 1) / E3_TILE_Y);
	E3ImageState::carve(nullptr, (size_t)width * height, T, &end);
	return (size_t)end + 256;
}

size_t e3_binning_bytes(int num_rendered)
{
	char* end = nullptr;
	E3BinState::carve(nullptr, (size_t)num_rendered, &end);
	return (size_t)end + 256;
}

// Caller-owned workspace of the no-sync path (e3_rasterize_forward_raw_ws): three buffers sized by
// the caller, the binning one for `cap` instances; R is copied to `r_host` asynchronously.
struct WsMode { char* g; char* b; size_t cap; char* i; int* r_host; };

static int forward_impl(
	const WsMode* ws,
	e3_alloc_fn geometry_alloc, void* geometry_ctx,
	e3_alloc_fn binning_alloc, void* binning_ctx,
	e3_alloc_fn image_alloc, void* image_ctx,
	int P, int D, int M, int channels,
	const float* background, int width, int height,
	const float* means3D, const float* shs, const float* f_rest, bool raw, const float* colors_precomp,
	const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, int* radii, int debug, void* stream, int* num_rendered_host)
{
	cudaStream_t st = (cudaStream_t)stream;
	if (num_rendered_host) *num_rendered_host = 0;
	if (P < 0 || width <= 0 || height <= 0 || (channels != 1 && channels != 3)) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if ((!ws && (!geometry_alloc || !binning_alloc || !image_alloc || !num_rendered_host)) || !means3D || !opacities || !background ||
		!viewmatrix || !projmatrix || !out_color || !radii)
		return E3_ERR_BAD_ARGUMENT;
	if (ws && (!ws->g || !ws->i || (ws->cap > 0 && !ws->b) || !ws->r_host || ws->cap > 0x7fffffffu)) return E3_ERR_BAD_ARGUMENT;
	if (!colors_precomp && (!shs || !cam_pos)) return E3_ERR_BAD_ARGUMENT;
	if (!colors_precomp && channels != 3 && !raw) return E3_ERR_UNSUPPORTED;
	if (!cov3D_precomp && (!scales || !rotations)) return E3_ERR_BAD_ARGUMENT;

	E3Frame f = make_frame(P, D, M, channels, width, height, tan_fovx, tan_fovy, scale_modifier);
	const size_t N = (size_t)width * height, T = (size_t)f.gx * f.gy;

	char* gbuf = ws ? ws->g : geometry_alloc(geometry_ctx, e3_geometry_bytes(P));
	if (!gbuf) return E3_ERR_ALLOC_FAILED;
	E3GeomState g = E3GeomState::carve(gbuf, P);
	char* ibuf = ws ? ws->i : image_alloc(image_ctx, e3_image_bytes(width, height));
	if (!ibuf) return E3_ERR_ALLOC_FAILED;
	E3ImageState im = E3ImageState::carve(ibuf, N, T);

	E3_CHECK(cudaMemsetAsync(g.flags, 0, 4 * sizeof(int), st));
	E3_STAGE(E3_STAGE_PREPROCESS_FWD, e3_launch_preprocess(f, means3D, scales, rotations, opacities, shs, f_rest, raw, cov3D_precomp, colors_precomp,
		viewmatrix, projmatrix, cam_pos, prefiltered != 0, radii, g, background, st));

	// depth order of the Gaussians (culled ones carry key 0xffffffff and emit nothing)
	int dres = 0;
	// culled Gaussians carry the key 0xffffffff and no instances: where they end up in the order is irrelevant
	E3_STAGE(E3_STAGE_DEPTH_SORT, e3_radix_sort_pairs(g.dkey, g.dval, (size_t)P, 32, g.sort_counts, g.scan_tmp, 0, &dres, st, nullptr, true));
	const uint32_t* order = g.dval[dres];
	// instance offsets in depth order; grand total = num_rendered
	E3_STAGE(E3_STAGE_OFFSETS_SCAN, e3_exclusive_scan_u32(g.tiles_touched, order, g.offsets, (size_t)P, g.scan_tmp,
		reinterpret_cast<uint32_t*>(g.flags), st));

	int R;
	char* bbuf;
	const uint32_t* n_dev = nullptr;
	if (ws) {
		// no host round trip: R stays on the device (flags[0]); everything downstream is launched for
		// the caller's capacity and bounds itself by min(R, cap). The host copy of {R, prefiltered flag}
		// lands in r_host[0..1] when the stream gets there.
		E3_CHECK(cudaMemcpyAsync(ws->r_host, g.flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
		R = (int)ws->cap;
		bbuf = ws->b;
		n_dev = reinterpret_cast<const uint32_t*>(g.flags);
	} else {
		int h_flags[2] = { 0, 0 };
		E3_CHECK(cudaMemcpyAsync(h_flags, g.flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
		E3_CHECK(cudaStreamSynchronize(st));        // the reference's blocking read of R (rasterizer_impl.cu:281)
		if (h_flags[1]) return E3_ERR_PREFILTERED;
		R = h_flags[0];
		*num_rendered_host = R;
		bbuf = binning_alloc(binning_ctx, e3_binning_bytes(R));
		if (!bbuf) return E3_ERR_ALLOC_FAILED;
	}
	E3BinState b = E3BinState::carve(bbuf, (size_t)R);

	// emit into the ping-pong slot from which the tile sort ends in slot 0, where backward looks
	const int tile_bits = (int)higher_msb((uint32_t)T);
	const int s0 = (R > 0) ? (e3_radix_passes(tile_bits) & 1) : 0;
	E3_STAGE(E3_STAGE_EMIT, e3_launch_emit_instances(P, order, g.offsets, g.tiles_touched, g.rect, f.gx, b.tkey[s0], b.val[s0], (uint32_t)R, st));
	int bres = 0;
	E3_STAGE(E3_STAGE_TILE_SORT, e3_radix_sort_pairs(b.tkey, b.val, (size_t)R, tile_bits, b.sort_counts, b.scan_tmp, s0, &bres, st, n_dev));
	if (bres != 0) return E3_ERR_BAD_ARGUMENT;
	E3_STAGE(E3_STAGE_TILE_RANGES, e3_launch_tile_ranges((size_t)R, b.tkey[0], im.ranges, T, st, n_dev));
	E3_STAGE(E3_STAGE_TILE_RANGES, e3_launch_tile_order(im.ranges, (int)T, im.tile_order, st));
	// single-channel raw mode blends against the luma of the background (written by the preprocess kernel)
	const float* bg_used = (raw && channels == 1) ? reinterpret_cast<const float*>(g.flags) + 2 : background;
	E3_STAGE(E3_STAGE_RENDER_FWD, e3_launch_render_forward(f, im.ranges, im.tile_order, b.val[0], g.rec, bg_used, im.final_T, im.n_contrib, out_color, st,
		/* exact exp: the drop-in path; the fused (raw-parameter) path evaluates what its backward evaluates */ !raw));
	return 0;
}

int e3_rasterize_forward(
	e3_alloc_fn geometry_alloc, void* geometry_ctx,
	e3_alloc_fn binning_alloc, void* binning_ctx,
	e3_alloc_fn image_alloc, void* image_ctx,
	int P, int D, int M, int channels,
	const float* background, int width, int height,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, int* radii, int debug, void* stream, int* num_rendered_host)
{
	return forward_impl(nullptr, geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, P, D, M, channels,
		background, width, height, means3D, shs, nullptr, false, colors_precomp, opacities, scales, scale_modifier, rotations,
		cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color, radii, debug, stream,
		num_rendered_host);
}

int e3_rasterize_forward_raw(
	e3_alloc_fn geometry_alloc, void* geometry_ctx,
	e3_alloc_fn binning_alloc, void* binning_ctx,
	e3_alloc_fn image_alloc, void* image_ctx,
	int P, int active_sh_degree, int channels,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy,
	float* out_color, int* radii, int debug, void* stream, int* num_rendered_host)
{
	if (active_sh_degree < 0 || active_sh_degree > 3 || (channels != 1 && channels != 3)) return E3_ERR_BAD_ARGUMENT;
	if (P > 0 && (!features_dc || (active_sh_degree > 0 && !features_rest))) return E3_ERR_BAD_ARGUMENT;
	return forward_impl(nullptr, geometry_alloc, geometry_ctx, binning_alloc, binning_ctx, image_alloc, image_ctx, P, active_sh_degree, 16, channels,
		background, width, height, xyz, features_dc, features_rest, true, nullptr, opacity_logits, log_scales, scale_modifier,
		raw_rotations, nullptr, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, 0, out_color, radii, debug, stream,
		num_rendered_host);
}

int e3_rasterize_forward_raw_ws(
	char* geometry_buffer, char* binning_buffer, size_t binning_capacity, char* image_buffer, int* num_rendered_host,
	int P, int active_sh_degree, int channels,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy,
	float* out_color, int* radii, int debug, void* stream)
{
	if (active_sh_degree < 0 || active_sh_degree > 3 || (channels != 1 && channels != 3)) return E3_ERR_BAD_ARGUMENT;
	if (P > 0 && (!features_dc || (active_sh_degree > 0 && !features_rest))) return E3_ERR_BAD_ARGUMENT;
	WsMode ws = { geometry_buffer, binning_buffer, binning_capacity, image_buffer, num_rendered_host };
	return forward_impl(&ws, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, P, active_sh_degree, 16, channels,
		background, width, height, xyz, features_dc, features_rest, true, nullptr, opacity_logits, log_scales, scale_modifier,
		raw_rotations, nullptr, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, 0, out_color, radii, debug, stream, nullptr);
}

int e3_rasterize_forward_ws(
	char* geometry_buffer, char* binning_buffer, size_t binning_capacity, char* image_buffer, int* num_rendered_host,
	int P, int D, int M, int channels,
	const float* background, int width, int height,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, int* radii, int debug, void* stream)
{
	if (P > 0 && (!geometry_buffer || !binning_buffer || !image_buffer || !num_rendered_host || binning_capacity == 0))
		return E3_ERR_BAD_ARGUMENT;
	WsMode ws = { geometry_buffer, binning_buffer, binning_capacity, image_buffer, num_rendered_host };
	return forward_impl(&ws, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, P, D, M, channels,
		background, width, height, means3D, shs, nullptr, false, colors_precomp, opacities, scales, scale_modifier, rotations,
		cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered, out_color, radii, debug, stream, nullptr);
}

int e3_rasterize_backward(
	int P, int D, int M, int R, int channels,
	const float* background, int width, int height,
	const float* means3D, const float* shs, const float* colors_precomp,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, const int* radii,
	char* geometry_buffer, char* binning_buffer, char* image_buffer,
	const float* dL_dpix, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
	float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
	int debug, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	if (P < 0 || (channels != 1 && channels != 3)) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (!geometry_buffer || !image_buffer || (R > 0 && !binning_buffer) || !radii || !means3D || !dL_dpix ||
		!dL_dmean2D || !dL_dopacity || !dL_dcolor || !dL_dmean3D || !dL_dcov3D || !viewmatrix || !projmatrix)
		return E3_ERR_BAD_ARGUMENT;
	if (shs && (!dL_dsh || !campos)) return E3_ERR_BAD_ARGUMENT;
	if (scales && (!rotations || !dL_dscale || !dL_drot)) return E3_ERR_BAD_ARGUMENT;
	(void)colors_precomp;   // colours live in the packed records written by the forward

	E3Frame f = make_frame(P, D, M, channels, width, height, tan_fovx, tan_fovy, scale_modifier);
	const size_t N = (size_t)width * height, T = (size_t)f.gx * f.gy;
	E3GeomState g = E3GeomState::carve(geometry_buffer, P);
	E3ImageState im = E3ImageState::carve(image_buffer, N, T);
	if (R > 0) {
		E3BinState b = E3BinState::carve(binning_buffer, (size_t)R);
		E3_STAGE(E3_STAGE_RENDER_BWD, e3_launch_render_backward(f, im.ranges, im.tile_order, b.val[0], g.rec, background, im.final_T, im.n_contrib,
			dL_dpix, g.acc, st, nullptr, /* the forward's exact exp: see render_bwd_kernel */ true));
	}
	E3_STAGE(E3_STAGE_PREPROCESS_BWD, e3_launch_preprocess_backward(f, means3D, radii, shs, nullptr, nullptr, false, false, scales, rotations,
		cov3D_precomp, viewmatrix, projmatrix, campos, g, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dmean3D, dL_dcov3D, dL_dsh,
		nullptr, dL_dscale, dL_drot, st));
	return 0;
}

// phase bit 0: backward blend (into the packed accumulators); bit 1: backward preprocess (to the outputs)
static int backward_raw_impl(int phase,
	int P, int active_sh_degree, int channels, int R,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, const int* radii,
	char* geometry_buffer, char* binning_buffer, char* image_buffer,
	const float* dL_dpix, int accumulate,
	float* d_xyz, float* d_features_dc, float* d_features_rest, float* d_opacity_logits,
	float* d_log_scales, float* d_raw_rotations, float* d_mean2D,
	int debug, void* stream, long long mc_delta = 0, const float* dL_dpix_stats = nullptr)
{
	cudaStream_t st = (cudaStream_t)stream;
	if (P < 0 || active_sh_degree < 0 || active_sh_degree > 3 || (channels != 1 && channels != 3)) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (mc_delta != 0 && !accumulate) return E3_ERR_BAD_ARGUMENT;      // multicast reductions only accumulate
	if (!geometry_buffer || !image_buffer || (R > 0 && !binning_buffer) || !radii || !xyz || !features_dc || !opacity_logits ||
		!log_scales || !raw_rotations || !dL_dpix || !d_xyz || !d_features_dc || !d_opacity_logits || !d_log_scales ||
		!d_raw_rotations || !viewmatrix || !projmatrix || !campos)
		return E3_ERR_BAD_ARGUMENT;
	if (active_sh_degree > 0 && (!features_rest || !d_features_rest)) return E3_ERR_BAD_ARGUMENT;
	E3Frame f = make_frame(P, active_sh_degree, 16, channels, width, height, tan_fovx, tan_fovy, scale_modifier);
	f.mc_delta = mc_delta;
	if (dL_dpix_stats && (channels != 3 || !d_mean2D)) return E3_ERR_BAD_ARGUMENT;
	f.mean2d_from_second_seed = dL_dpix_stats != nullptr;
	const size_t N = (size_t)width * height, T = (size_t)f.gx * f.gy;
	E3GeomState g = E3GeomState::carve(geometry_buffer, P);
	E3ImageState im = E3ImageState::carve(image_buffer, N, T);
	const float* bg_used = channels == 1 ? reinterpret_cast<const float*>(g.flags) + 2 : background;
	if (R > 0 && (phase & 1)) {
		E3BinState b = E3BinState::carve(binning_buffer, (size_t)R);
		E3_STAGE(E3_STAGE_RENDER_BWD, e3_launch_render_backward(f, im.ranges, im.tile_order, b.val[0], g.rec, bg_used, im.final_T, im.n_contrib,
			dL_dpix, g.acc, st, dL_dpix_stats));
	}
	if (phase & 2)
		E3_STAGE(E3_STAGE_PREPROCESS_BWD, e3_launch_preprocess_backward(f, xyz, radii, features_dc, features_rest, opacity_logits, true,
			accumulate != 0, log_scales, raw_rotations, nullptr, viewmatrix, projmatrix, campos, g, d_mean2D, nullptr, d_opacity_logits,
			nullptr, d_xyz, nullptr, d_features_dc, active_sh_degree > 0 ? d_features_rest : nullptr, d_log_scales, d_raw_rotations, st));
	return 0;
}

int e3_rasterize_backward_raw(
	int P, int active_sh_degree, int channels, int R,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, const int* radii,
	char* geometry_buffer, char* binning_buffer, char* image_buffer,
	const float* dL_dpix, int accumulate,
	float* d_xyz, float* d_features_dc, float* d_features_rest, float* d_opacity_logits,
	float* d_log_scales, float* d_raw_rotations, float* d_mean2D,
	int debug, void* stream)
{
	return backward_raw_impl(3, P, active_sh_degree, channels, R, background, width, height, xyz, features_dc, features_rest,
		opacity_logits, log_scales, scale_modifier, raw_rotations, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
		geometry_buffer, binning_buffer, image_buffer, dL_dpix, accumulate, d_xyz, d_features_dc, d_features_rest, d_opacity_logits,
		d_log_scales, d_raw_rotations, d_mean2D, debug, stream);
}

int e3_rasterize_backward_raw_dual(
	int P, int active_sh_degree, int channels, int R,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* campos,
	float tan_fovx, float tan_fovy, const int* radii,
	char* geometry_buffer, char* binning_buffer, char* image_buffer,
	const float* dL_dpix, const float* dL_dpix_stats, int accumulate,
	float* d_xyz, float* d_features_dc, float* d_features_rest, float* d_opacity_logits,
	float* d_log_scales, float* d_raw_rotations, float* d_mean2D,
	int debug, void* stream)
{
	return backward_raw_impl(3, P, active_sh_degree, channels, R, background, width, height, xyz, features_dc, features_rest,
		opacity_logits, log_scales, scale_modifier, raw_rotations, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, radii,
		geometry_buffer, binning_buffer, image_buffer, dL_dpix, accumulate, d_xyz, d_features_dc, d_features_rest, d_opacity_logits,
		d_log_scales, d_raw_rotations, d_mean2D, debug, stream, 0, dL_dpix_stats);
}

int e3_event_loss(int width, int height, const float* image, const float* image_now, const float* image_next,
	const void* gt_now, const void* gt_next, const void* gt_intensity, const void* gt_blur, int gt_is_u8, const float* c,
	float upstream, int event_channels, void* scratch, size_t scratch_bytes, float* d_image, float* d_image_now,
	float* d_image_next, float* out5, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	const int debug = 0;
	if ((event_channels != 1 && event_channels != 3 && event_channels != E3_EVENT_SHARED_FIRST_VIEW) || width <= 0 || height <= 0 || !image || !image_now || !image_next || !gt_now || !gt_next || !gt_intensity || !c ||
		!scratch || scratch_bytes < 8 * sizeof(double) || (reinterpret_cast<size_t>(scratch) & 7) || !d_image || !d_image_now ||
		!d_image_next || !out5)
		return E3_ERR_BAD_ARGUMENT;
	E3_STAGE(E3_STAGE_EVENT_LOSS, e3_launch_event_gt_fraction((size_t)width * height, gt_now, gt_next, gt_is_u8, static_cast<double*>(scratch), st));
	E3_STAGE(E3_STAGE_EVENT_LOSS, e3_launch_event_loss((size_t)width * height, image, image_now, image_next, gt_now, gt_next,
		gt_intensity, gt_blur, gt_is_u8, c, upstream, event_channels, static_cast<double*>(scratch), d_image, d_image_now, d_image_next, out5, st));
	return 0;
}

int e3_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
	uint8_t* present, void* stream)
{
	(void)projmatrix;   // the reference's test only uses the view-space depth (auxiliary.h:154)
	if (P < 0) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (!means3D || !viewmatrix || !present) return E3_ERR_BAD_ARGUMENT;
	E3_CHECK(e3_launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream));
	return 0;
}

size_t e3_scan_scratch_bytes(size_t n) { return (e3_scan_blocks(n) + 2) * sizeof(uint32_t); }

int e3_exclusive_scan(const uint32_t* in, uint32_t* out, size_t n, void* scratch, size_t scratch_bytes,
	uint32_t* total, void* stream)
{
	if (n == 0) {
		if (total) E3_CHECK(cudaMemsetAsync(total, 0, sizeof(uint32_t), (cudaStream_t)stream));
		return 0;
	}
	if (!in || !out || !scratch || scratch_bytes < e3_scan_scratch_bytes(n)) return E3_ERR_BAD_ARGUMENT;
	E3_CHECK(e3_exclusive_scan_u32(in, nullptr, out, n, static_cast<uint32_t*>(scratch), total, (cudaStream_t)stream));
	return 0;
}

// Two side streams and their fork/join events per device: the three views of an iteration are
// independent until the loss and again until the gradient accumulation, and at event-camera
// resolutions one view alone does not fill the 148 SMs (374 tiles at 346x260).
// aux: what needs no render (zeroing the gradient rows, the GT-only part of the loss) runs beside the forwards
struct ViewStreams { cudaStream_t side[2], aux; cudaEvent_t fork, join[2], aux_join; bool ok; };
static ViewStreams g_view_streams[64];
static std::mutex g_view_streams_mu;

static cudaError_t view_streams(ViewStreams** out)
{
	int dev = 0;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return e;
	if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
	std::lock_guard<std::mutex> l(g_view_streams_mu);
	ViewStreams& v = g_view_streams[dev];
	if (!v.ok) {
		for (int i = 0; i < 2; i++) {
			if ((e = cudaStreamCreateWithFlags(&v.side[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
			if ((e = cudaEventCreateWithFlags(&v.join[i], cudaEventDisableTiming)) != cudaSuccess) return e;
		}
		if ((e = cudaEventCreateWithFlags(&v.fork, cudaEventDisableTiming)) != cudaSuccess) return e;
		if ((e = cudaStreamCreateWithFlags(&v.aux, cudaStreamNonBlocking)) != cudaSuccess) return e;
		if ((e = cudaEventCreateWithFlags(&v.aux_join, cudaEventDisableTiming)) != cudaSuccess) return e;
		v.ok = true;
	}
	*out = &v;
	return cudaSuccess;
}

int e3_event_iteration(const e3_event_iteration_args* a, void* stream)
{
	if (!a) return E3_ERR_BAD_ARGUMENT;
	// views[1].channels == 0: the first event view IS the intensity view (same camera) and is rendered once
	const bool shared = a->views[1].channels == 0;
	if (a->views[0].channels != 3) return E3_ERR_BAD_ARGUMENT;
	if (shared ? (a->views[2].channels != 1 || !a->views[1].seed || !a->d_mean2D) : (a->views[1].channels != a->views[2].channels))
		return E3_ERR_BAD_ARGUMENT;
	cudaStream_t main_st = (cudaStream_t)stream;
	ViewStreams* vs = nullptr;
	E3_CHECK(view_streams(&vs));
	cudaStream_t sts[3] = { main_st, vs->side[0], vs->side[1] };
	// Whatever way this call ends, the caller's stream is ordered behind everything the side streams were given:
	// after an error return the caller may free or reuse the workspace, seed and gradient buffers on its stream.
	struct JoinSideStreams {
		ViewStreams* v; cudaStream_t m; bool armed = true;
		~JoinSideStreams()
		{
			if (!armed) return;
			for (int k = 0; k < 2; k++)
				if (cudaEventRecord(v->join[k], v->side[k]) == cudaSuccess) cudaStreamWaitEvent(m, v->join[k], 0);
			if (cudaEventRecord(v->aux_join, v->aux) == cudaSuccess) cudaStreamWaitEvent(m, v->aux_join, 0);
		}
	} join_guard{ vs, main_st };
	if (a->grad_multicast_delta != 0 && !a->accumulate) return E3_ERR_BAD_ARGUMENT;   // the caller zeroes and fences (see header)
	if (a->P > 0 && (!a->gt_now || !a->gt_next || !a->gt_intensity || !a->c || !a->loss_scratch || (reinterpret_cast<size_t>(a->loss_scratch) & 7) ||
		!a->out5 || !a->views[0].seed || !a->views[1].seed || !a->views[2].seed))
		return E3_ERR_BAD_ARGUMENT;
	// ---- three forwards, one stream each; beside them what needs no render: the gradient rows are zeroed and the
	// GT-only part of the loss is taken, off the critical path between the forwards and the backward blends ----
	E3_CHECK(cudaEventRecord(vs->fork, main_st));
	for (int k = 1; k < 3; k++) E3_CHECK(cudaStreamWaitEvent(sts[k], vs->fork, 0));
	if (a->P > 0) {
		const int debug = 0;
		cudaStream_t st = vs->aux;
		E3_CHECK(cudaStreamWaitEvent(st, vs->fork, 0));
		if (a->wait_before_loss_event) E3_CHECK(cudaStreamWaitEvent(st, (cudaEvent_t)a->wait_before_loss_event, 0));
		E3_STAGE(E3_STAGE_EVENT_LOSS, e3_launch_event_gt_fraction((size_t)a->width * a->height, a->gt_now, a->gt_next, a->gt_is_u8,
			static_cast<double*>(a->loss_scratch), st));
		if (!a->accumulate) {
			const size_t P = (size_t)a->P;
			E3_CHECK(cudaMemsetAsync(a->d_xyz, 0, P * 3 * sizeof(float), st));
			E3_CHECK(cudaMemsetAsync(a->d_features_dc, 0, P * 3 * sizeof(float), st));
			if (a->d_features_rest && a->active_sh_degree > 0) E3_CHECK(cudaMemsetAsync(a->d_features_rest, 0, P * 45 * sizeof(float), st));
			E3_CHECK(cudaMemsetAsync(a->d_opacity_logits, 0, P * sizeof(float), st));
			E3_CHECK(cudaMemsetAsync(a->d_log_scales, 0, P * 3 * sizeof(float), st));
			E3_CHECK(cudaMemsetAsync(a->d_raw_rotations, 0, P * 4 * sizeof(float), st));
		}
		E3_CHECK(cudaEventRecord(vs->aux_join, st));
	}
	for (int k = 0; k < 3; k++) {
		if (shared && k == 1) continue;
		const e3_view& v = a->views[k];
		int rc = e3_rasterize_forward_raw_ws(v.geometry_buffer, v.binning_buffer, (size_t)a->binning_capacity, v.image_buffer,
			v.num_rendered_host, a->P, a->active_sh_degree, v.channels, a->background, a->width, a->height, a->xyz, a->features_dc,
			a->features_rest, a->opacity_logits, a->log_scales, a->scale_modifier, a->raw_rotations, v.viewmatrix, v.projmatrix,
			v.cam_pos, v.tan_fovx, v.tan_fovy, v.out_color, v.radii, 0, sts[k]);
		if (rc) return rc;
	}
	for (int k = 1; k < 3; k++) {
		E3_CHECK(cudaEventRecord(vs->join[k - 1], sts[k]));
		E3_CHECK(cudaStreamWaitEvent(main_st, vs->join[k - 1], 0));
	}
	if (a->forwards_done_event) E3_CHECK(cudaEventRecord((cudaEvent_t)a->forwards_done_event, main_st));
	if (a->P == 0) return 0;
	if (a->wait_before_loss_event) E3_CHECK(cudaStreamWaitEvent(main_st, (cudaEvent_t)a->wait_before_loss_event, 0));
	E3_CHECK(cudaStreamWaitEvent(main_st, vs->aux_join, 0));
	{
		const int debug = 0;
		cudaStream_t st = main_st;
		E3_STAGE(E3_STAGE_EVENT_LOSS, e3_launch_event_loss((size_t)a->width * a->height, a->views[0].out_color,
			a->views[shared ? 0 : 1].out_color, a->views[2].out_color, a->gt_now, a->gt_next, a->gt_intensity, a->gt_blur, a->gt_is_u8, a->c,
			a->loss_upstream != 0.f ? a->loss_upstream : 1.0f, shared ? E3_EVENT_SHARED_FIRST_VIEW : a->views[1].channels,
			static_cast<double*>(a->loss_scratch), a->views[0].seed, a->views[1].seed, a->views[2].seed, a->out5, st));
	}
	int rc = 0;
	// ---- per view, on its own stream: backward blend, then the gradient accumulation. The three
	// views add into the same (zeroed) gradient tensors with vector reductions, so nothing serialises
	// behind the blends and the latency-bound accumulation kernels fill the gaps the blends leave. ----
	E3_CHECK(cudaEventRecord(vs->fork, main_st));
	for (int k = 1; k < 3; k++) E3_CHECK(cudaStreamWaitEvent(sts[k], vs->fork, 0));
	for (int k = 0; k < 3; k++) {
		if (shared && k == 1) continue;
		const e3_view& v = a->views[k];
		{
			// shared first view: one backward blend with two seeds -- views[1].seed holds the total seed of the
			// view (parameter gradients), views[0].seed the intensity term's (d_mean2D), see e3_rasterize_backward_raw_dual
			const float* seed = (shared && k == 0) ? a->views[1].seed : v.seed;
			const float* stats = (shared && k == 0) ? a->views[0].seed : nullptr;
			rc = backward_raw_impl(3, a->P, a->active_sh_degree, v.channels, (int)a->binning_capacity, a->background, a->width,
				a->height, a->xyz, a->features_dc, a->features_rest, a->opacity_logits, a->log_scales, a->scale_modifier,
				a->raw_rotations, v.viewmatrix, v.projmatrix, v.cam_pos, v.tan_fovx, v.tan_fovy, v.radii, v.geometry_buffer,
				v.binning_buffer, v.image_buffer, seed, 1, a->d_xyz, a->d_features_dc,
				a->d_features_rest, a->d_opacity_logits, a->d_log_scales, a->d_raw_rotations, k == 0 ? a->d_mean2D : nullptr, 0,
				sts[k], a->grad_multicast_delta, stats);
			if (rc) return rc;
		}
	}
	for (int k = 1; k < 3; k++) {
		E3_CHECK(cudaEventRecord(vs->join[k - 1], sts[k]));
		E3_CHECK(cudaStreamWaitEvent(main_st, vs->join[k - 1], 0));
	}
	join_guard.armed = false;
	return 0;
}

size_t e3_l1_ssim_scratch_bytes(int width, int height, int gray)
{
	return (width > 0 && height > 0) ? e3_l1_ssim_scratch_words(width, height, gray != 0) * sizeof(float) : 0;
}

int e3_l1_ssim_loss(int width, int height, int gray, const float* image, const void* gt, int gt_is_u8, float lambda_dssim,
	float upstream, void* scratch, size_t scratch_bytes, float* d_image, float* out3, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	const int debug = 0;
	if (width <= 0 || height <= 0 || !image || !gt || !scratch || !d_image || !out3 || (reinterpret_cast<size_t>(scratch) & 7) ||
		scratch_bytes < e3_l1_ssim_scratch_bytes(width, height, gray))
		return E3_ERR_BAD_ARGUMENT;
	E3_STAGE(E3_STAGE_EVENT_LOSS, e3_launch_l1_ssim(width, height, gray != 0, image, gt, gt_is_u8 != 0, lambda_dssim, upstream,
		static_cast<float*>(scratch), d_image, out3, st));
	return 0;
}

size_t e3_dist2_scratch_bytes(int P)
{
	return P > 0 ? e3_knn_scratch_layout(P, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr) : 0;
}

int e3_dist2(int P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes, void* stream)
{
	cudaStream_t st = (cudaStream_t)stream;
	const int debug = 0;
	if (P < 0) return E3_ERR_BAD_ARGUMENT;
	if (P == 0) return 0;
	if (!points || !mean_dist2 || !scratch || scratch_bytes < e3_dist2_scratch_bytes(P)) return E3_ERR_BAD_ARGUMENT;
	E3_STAGE(E3_STAGE_DIST2, e3_launch_dist2(P, points, mean_dist2, static_cast<char*>(scratch), st));
	return 0;
}

unsigned long long e3_kernel_launches(void) { return g_e3_launches.load(); }

int e3_profile_enable(unsigned stage_mask)
{
	std::lock_guard<std::mutex> l(g_prof_mu);
	g_prof_mask = stage_mask;
	return 0;
}

int e3_profile_read(double* ms_sum, long long* count, int reset)
{
	std::lock_guard<std::mutex> l(g_prof_mu);
	for (const ProfSpan& sp : g_prof_open) {
		cudaError_t e = cudaEventSynchronize(sp.b);
		if (e != cudaSuccess) return (int)e;
		float ms = 0.f;
		e = cudaEventElapsedTime(&ms, sp.a, sp.b);
		if (e != cudaSuccess) return (int)e;
		g_prof_ms[sp.stage] += ms;
		g_prof_n[sp.stage] += 1;
		g_prof_pool.push_back(sp.a);
		g_prof_pool.push_back(sp.b);
	}
	g_prof_open.clear();
	for (int i = 0; i < E3_NUM_STAGES; i++) {
		if (ms_sum) ms_sum[i] = g_prof_ms[i];
		if (count) count[i] = g_prof_n[i];
		if (reset) { g_prof_ms[i] = 0.0; g_prof_n[i] = 0; }
	}
	return 0;
}

void* e3_geometry_ptr(char* geometry_buffer, int P, int which)
{
	E3GeomState g = E3GeomState::carve(geometry_buffer, (size_t)P);
	switch (which) {
	case 0: return g.depths;
	case 1: return g.rec;
	case 2: return g.cov3D;
	case 3: return g.tiles_touched;
	case 4: return g.dval[0];   // 4 passes of the depth sort end in slot 0
	case 5: return g.rgb;
	case 6: return g.clamped;
	}
	return nullptr;
}

void* e3_binning_ptr(char* binning_buffer, int num_rendered, int which)
{
	E3BinState b = E3BinState::carve(binning_buffer, (size_t)num_rendered);
	switch (which) {
	case 0: return b.val[0];
	case 1: return b.tkey[0];
	}
	return nullptr;
}

void* e3_image_ptr(char* image_buffer, int width, int height, int which)
{
	const size_t T = (size_t)((width + E3_TILE_X - 1) / E3_TILE_X) * ((height + E3_TILE_Y - 1) / E3_TILE_Y);
	E3ImageState im = E3ImageState::carve(image_buffer, (size_t)width * height, T);
	switch (which) {
	case 0: return im.final_T;
	case 1: return im.n_contrib;
	case 2: return im.ranges;
	}
	return nullptr;
}

}  // extern "C"
