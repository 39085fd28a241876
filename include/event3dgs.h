/*
 * event3dgs.h -- C ABI of libevent3dgs.so, the B200 (sm_100a) implementation of the
 * Event-3DGS differentiable splatting hot path.
 *
 * Plain pointers and sizes only; no torch, pybind or C++ types. Every pointer is a DEVICE
 * pointer unless its name ends in _host. All kernels are enqueued on `stream`
 * (a cudaStream_t passed as void*; NULL = the legacy default stream the reference uses).
 * The library allocates no device memory of its own: workspaces come from the caller through
 * the e3_alloc_fn callbacks (forward) or are handed back as pointers (backward), exactly like
 * the reference's resize-callback scheme. Return value: 0 on success, otherwise a cudaError_t
 * value (> 0) or an E3_ERR_* code (< 0); e3_error_string() explains both.
 *
 * Reference citations are relative to /root/reference/submodules/diff-gaussian-rasterization/.
 */
#ifndef EVENT3DGS_H_INCLUDED
#define EVENT3DGS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define E3_API __attribute__((visibility("default")))
#else
#define E3_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define E3_ERR_BAD_ARGUMENT  (-1)   /* shape / pointer contract violated */
#define E3_ERR_ALLOC_FAILED  (-2)   /* an e3_alloc_fn returned NULL */
#define E3_ERR_PREFILTERED   (-3)   /* a point was culled although `prefiltered` was set
                                       (the reference __trap()s, cuda_rasterizer/auxiliary.h:156-160) */
#define E3_ERR_UNSUPPORTED   (-4)   /* e.g. SH colours requested with channels != 3
                                       (cuda_rasterizer/rasterizer_impl.cu:242-245) */

E3_API const char* e3_error_string(int code);
E3_API int e3_version(void);

/* Replaces std::function<char*(size_t)> of CudaRasterizer::Rasterizer::forward
 * (cuda_rasterizer/rasterizer.h:32-34) and resizeFunctional (rasterize_points.cu:27-33):
 * must return a device buffer of at least `bytes` bytes, 256-byte aligned, that stays alive
 * until the matching backward has run. */
typedef char* (*e3_alloc_fn)(void* ctx, size_t bytes);

/* Sizes of the three opaque state buffers (replace required<GeometryState>(P),
 * required<ImageState>(N), required<BinningState>(R), cuda_rasterizer/rasterizer_impl.h:67-73).
 * The layouts are private to this library and differ from the reference's. */
E3_API size_t e3_geometry_bytes(int P);
E3_API size_t e3_image_bytes(int width, int height);
E3_API size_t e3_binning_bytes(int num_rendered);

/* Replaces CudaRasterizer::Rasterizer::forward (cuda_rasterizer/rasterizer.h:31-54,
 * rasterizer_impl.cu:198-336) as called by RasterizeGaussiansCUDA (rasterize_points.cu:35-115).
 * Same arguments in the same order, plus: `channels` (3 = the reference's NUM_CHANNELS,
 * config.h:15; 1 = single-channel "gray" colours, only with colors_precomp), the three
 * callback contexts, `stream`. Unused optional inputs are NULL (the reference passes the data
 * pointer of an empty tensor). out_color is [channels,H,W]; radii is int32[P] and must not be
 * NULL. *num_rendered_host receives R (this call synchronises the stream once to read R, as the
 * reference does at rasterizer_impl.cu:281). P == 0 is the caller's short-circuit
 * (rasterize_points.cu:81): the function then only returns 0 with R = 0. */
E3_API int e3_rasterize_forward(
	e3_alloc_fn geometry_alloc, void* geometry_ctx,
	e3_alloc_fn binning_alloc, void* binning_ctx,
	e3_alloc_fn image_alloc, void* image_ctx,
	int P, int D, int M, int channels,
	const float* background,
	int width, int height,
	const float* means3D,
	const float* shs,
	const float* colors_precomp,
	const float* opacities,
	const float* scales,
	float scale_modifier,
	const float* rotations,
	const float* cov3D_precomp,
	const float* viewmatrix,
	const float* projmatrix,
	const float* cam_pos,
	float tan_fovx, float tan_fovy,
	int prefiltered,
	float* out_color,
	int* radii,
	int debug,
	void* stream,
	int* num_rendered_host);

/* Replaces CudaRasterizer::Rasterizer::backward (cuda_rasterizer/rasterizer.h:56-84,
 * rasterizer_impl.cu:340-434) as called by RasterizeGaussiansBackwardCUDA
 * (rasterize_points.cu:117-196). Differences from the reference, all in the caller's favour:
 * the nine gradient arrays need NOT be zero-filled (every row is written, rows of culled
 * Gaussians with zeros), and dL_dconic may be NULL (it is an internal quantity there).
 * Shapes: dL_dpix [channels,H,W]; dL_dmean2D [P,3] (z = 0); dL_dopacity [P]; dL_dcolor
 * [P,channels]; dL_dmean3D [P,3]; dL_dcov3D [P,6]; dL_dsh [P,M,3]; dL_dscale [P,3]; dL_drot [P,4]. */
E3_API int e3_rasterize_backward(
	int P, int D, int M, int R, int channels,
	const float* background,
	int width, int height,
	const float* means3D,
	const float* shs,
	const float* colors_precomp,
	const float* scales,
	float scale_modifier,
	const float* rotations,
	const float* cov3D_precomp,
	const float* viewmatrix,
	const float* projmatrix,
	const float* campos,
	float tan_fovx, float tan_fovy,
	const int* radii,
	char* geometry_buffer,
	char* binning_buffer,
	char* image_buffer,
	const float* dL_dpix,
	float* dL_dmean2D,
	float* dL_dconic,
	float* dL_dopacity,
	float* dL_dcolor,
	float* dL_dmean3D,
	float* dL_dcov3D,
	float* dL_dsh,
	float* dL_dscale,
	float* dL_drot,
	int debug,
	void* stream);

/* Fused variants that take the LEAF parameters of the reference's GaussianModel
 * (scene/gaussian_model.py:124-147: xyz [P,3], features_dc [P,1,3], features_rest [P,15,3],
 * scaling = log-scales [P,3], rotation = unnormalised quaternions [P,4], opacity = logits [P,1]).
 * They replace, in one kernel each way, what render() does in ~60 PyTorch launches per call plus
 * their autograd backward: the activations exp / F.normalize / sigmoid
 * (scene/gaussian_model.py:95-115), torch.cat of the SH features (:109-113), the view direction and
 * eval_sh + clamp_min(x + 0.5, 0) (gaussian_renderer/__init__.py:74-81, utils/sh_utils.py:57-112),
 * followed by the rasterizer call. active_sh_degree in 0..3; features_rest may be NULL at degree 0.
 * channels = 3: RGB image [3,H,W]. channels = 1: the "--gray" rewrite for views that the loss only
 * consumes through rgb_to_LUVscale (utils/loss_utils.py:24-28): every Gaussian's clamped colour is
 * projected to Y = 0.4124 R + 0.35758 G + 0.1804 B before blending and the image is [1,H,W]; since
 * blending is linear in colour this equals the luma of the 3-channel render up to fp32 rounding.
 * `background` always points to 3 floats. */
E3_API int e3_rasterize_forward_raw(
	e3_alloc_fn geometry_alloc, void* geometry_ctx,
	e3_alloc_fn binning_alloc, void* binning_ctx,
	e3_alloc_fn image_alloc, void* image_ctx,
	int P, int active_sh_degree, int channels,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy,
	float* out_color, int* radii, int debug, void* stream, int* num_rendered_host);

/* e3_rasterize_forward without the host round trip (see e3_rasterize_forward_raw_ws for the contract:
 * caller-owned buffers, binning_capacity an estimate of num_rendered, {R, prefiltered flag} copied
 * asynchronously to num_rendered_host[0..1] in pinned memory, the caller verifies R <=
 * binning_capacity afterwards). Opt-in route of the drop-in binding (E3_DROPIN_NO_SYNC=1): the
 * reference's blocking read (cuda_rasterizer/rasterizer_impl.cu:281) is what keeps the host from
 * running ahead of the ~100 small ATen launches between two renders. For the backward pass hand
 * e3_rasterize_backward the same buffers and binning_capacity as its R. */
E3_API int e3_rasterize_forward_ws(
	char* geometry_buffer, char* binning_buffer, size_t binning_capacity, char* image_buffer, int* num_rendered_host,
	int P, int D, int M, int channels,
	const float* background, int width, int height,
	const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
	const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy, int prefiltered,
	float* out_color, int* radii, int debug, void* stream);

/* e3_rasterize_forward_raw without the host round trip. The reference blocks the host in every
 * forward to read num_rendered and size the binning buffer (cuda_rasterizer/rasterizer_impl.cu:
 * 281-285); at event-camera resolutions (346x260) that stall and the launch gaps around it cost more
 * than the kernels. Here the caller owns all three buffers for the whole training run:
 * geometry_buffer >= e3_geometry_bytes(P), image_buffer >= e3_image_bytes(W, H), binning_buffer >=
 * e3_binning_bytes(binning_capacity) with binning_capacity an estimate of num_rendered (e.g. 1.3x
 * the largest value seen). R stays on the device; every later kernel is launched for the capacity
 * and bounds itself by min(R, capacity). {R, prefiltered flag} are copied asynchronously to
 * num_rendered_host[0..1] (pinned host memory): once the stream has passed that copy the caller
 * checks R <= binning_capacity and, if not, enlarges the buffer and repeats the step (the results of
 * an overflowed call are defined but incomplete). For the backward pass hand
 * e3_rasterize_backward_raw the same buffers and binning_capacity as its R. */
E3_API int e3_rasterize_forward_raw_ws(
	char* geometry_buffer, char* binning_buffer, size_t binning_capacity, char* image_buffer,
	int* num_rendered_host,
	int P, int active_sh_degree, int channels,
	const float* background, int width, int height,
	const float* xyz, const float* features_dc, const float* features_rest, const float* opacity_logits,
	const float* log_scales, float scale_modifier, const float* raw_rotations,
	const float* viewmatrix, const float* projmatrix, const float* cam_pos,
	float tan_fovx, float tan_fovy,
	float* out_color, int* radii, int debug, void* stream);

/* Backward of e3_rasterize_forward_raw down to the leaf parameters (what loss.backward() produces
 * in the reference through _RasterizeGaussians.backward and the autograd graph of render()).
 * accumulate = 0: every row of the outputs is written (zeros for culled Gaussians);
 * accumulate = 1: visible rows are ADDED to the outputs (the three renders of an event iteration
 * sum into one gradient tensor, like autograd's accumulation into .grad). d_features_rest may be
 * NULL at degree 0 (it is identically zero). d_mean2D [P,3] is optional: the gradient of the
 * screen-space means that train.py:319-320 feeds to the densification statistics. */
E3_API int e3_rasterize_backward_raw(
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
	int debug, void* stream);

/* e3_rasterize_backward_raw with TWO seeds for one 3-channel render: all parameter gradients come from
 * dL_dpix, d_mean2D (required) comes from dL_dpix_stats alone. For an event iteration whose intensity
 * camera and first event camera are the same view (they are in the reference's datasets,
 * scene/dataset_readers.py:145-160): the reference renders that view twice (train.py:144,159-160);
 * rendered once, dL_dpix = seed of the intensity term + seed of the event term gives the same parameter
 * gradients (they are linear in the seed), and dL_dpix_stats = the intensity term's seed gives the
 * screen-space gradient that the reference's densification statistics take from its FIRST render
 * (train.py:145,319-320) -- so nothing observable changes. */
E3_API int e3_rasterize_backward_raw_dual(
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
	int debug, void* stream);

/* Event loss of train.py:165-203 (with utils/loss_utils.py:24-28, 234-249, 270-271), value and
 * backward seeds in one streaming pass over the renders behind a pass over the two GT event frames. Inputs are [3,H,W] float images: the three renders
 * (image = intensity view, image_now / image_next = event views index / index+1), the GT event
 * frames, the GT intensity frame and, for --deblur, the blurry frame (else NULL). The four GT
 * frames are float32 in [0,1] (gt_is_u8 = 0) or the raw 8-bit pixels (gt_is_u8 = 1; the kernel
 * applies the reference's /255, utils/general_utils.py:21-27). `c` points to the
 * learnable event threshold on the device (train.py:71,165); the GT threshold is the constant 0.17
 * (train.py:170). `upstream` scales all gradients (1 for loss.backward()). event_channels = 3:
 * image_now / image_next and their seeds are [3,H,W]; 1: they are single-channel luma renders
 * (channels = 1 of e3_rasterize_forward_raw) and the seeds are [1,H,W]. scratch: >= 64 bytes,
 * 8-byte aligned device memory. Outputs: the three seeds dL/dimage* [3,H,W] and
 * out5 = { loss, dL/dc, event L1, intensity L1, fraction of pixels with non-zero GT event }. */
/* event_channels value for a render triplet whose first event view IS the intensity view (rendered once):
 * image_now must be the 3-channel `image` itself, image_next a 1-channel luma render; d_image_now then
 * receives the TOTAL seed of the shared view (event term + intensity term) and d_image the intensity
 * term alone -- the two seeds of e3_rasterize_backward_raw_dual. */
#define E3_EVENT_SHARED_FIRST_VIEW 0x31
E3_API int e3_event_loss(int width, int height, const float* image, const float* image_now,
	const float* image_next, const void* gt_now, const void* gt_next, const void* gt_intensity,
	const void* gt_blur, int gt_is_u8, const float* c, float upstream, int event_channels, void* scratch,
	size_t scratch_bytes, float* d_image, float* d_image_now, float* d_image_next, float* out5,
	void* stream);

/* One Adam step over up to E3_ADAM_MAX_GROUPS parameter groups in a single kernel launch.
 * Replaces gaussians.optimizer.step() (train.py:330-332) on the optimizer that
 * scene/gaussian_model.py:149-175 builds as torch.optim.Adam(l, lr=0.0, eps=1e-15) with one group
 * per leaf tensor, and optimizer_c.step() (train.py:71-73,212). Same arithmetic as torch (no
 * amsgrad, no weight decay): `step` is the 1-based step count t used for the bias corrections.
 * groups_host is an array in HOST memory; the pointers inside are device pointers to n floats. */
#define E3_ADAM_MAX_GROUPS 8
typedef struct e3_adam_group {
	float* param;
	const float* grad;
	float* exp_avg;
	float* exp_avg_sq;
	unsigned long long n;
	float lr;
	float eps;
} e3_adam_group;
E3_API int e3_adam_step(const e3_adam_group* groups_host, int num_groups, int step, double beta1,
	double beta2, void* stream);

/* Sharded Adam fused with the gradient exchange of the replicated multi-GPU trainer (SURVEY.md 8e upgrade path;
 * no counterpart in the single-GPU reference, whose optimizer step is scene/gaussian_model.py:149-175 +
 * train.py:330-332). Groups as for e3_adam_step, but `param` and `grad` point into buffers that exist in symmetric
 * memory on all `world` ranks, with the byte offsets from them to their NVLS multicast aliases given; both are laid
 * out in segments padded to 16 bytes, the moments are ordinary local arrays of n floats. Rank r owns the r-th
 * 1/world of every group's 16-byte vectors: it fetches grad_scale * (sum over ranks of the gradient) from the switch
 * (multimem.ld_reduce), applies Adam to those rows and its rows of the moments, and writes the new parameters into
 * every rank's copy (multimem.st). One kernel. The caller places a cross-rank barrier on the stream before (all local
 * gradients complete) and after (all parameters written). Moments of rows a rank does not own are not touched. */
E3_API int e3_nvls_sharded_adam(const e3_adam_group* groups_host, int num_groups, int step, double beta1, double beta2,
	long long grad_multicast_delta, long long param_multicast_delta, int rank, int world, float grad_scale, void* stream);

/* Densification statistics of one iteration (train.py:317-320, scene/gaussian_model.py:405-407),
 * one launch: for every Gaussian with radii > 0 (the visibility filter of the first render)
 *   max_radii2D = max(max_radii2D, radii); xyz_gradient_accum += |d_mean2D[:, :2]|; denom += 1.
 * d_mean2D is the [P,3] screen-space gradient (viewspace_points.grad). */
E3_API int e3_densification_stats(int P, const int* radii, const float* d_mean2D, float* max_radii2D,
	float* xyz_gradient_accum, float* denom, void* stream);

/* Adaptive density control, replacing GaussianModel.densify_and_prune (scene/gaussian_model.py:
 * 389-403) together with the tensor / optimizer-state surgery beneath it (densify_and_clone
 * :374-387, densify_and_split :349-372, densification_postfix :329-347, cat_tensors_to_optimizer
 * :307-327, prune_points :291-305, _prune_optimizer :273-289). Two calls, because the new row count
 * is only known after the first and the caller owns all memory:
 *
 *  e3_densify_plan   evaluates the clone / split / prune criteria on the P current rows and builds
 *                    the row map in `scratch` (e3_densify_scratch_bytes(P) bytes of device memory).
 *                    counts (device, 8 ints): [0] kept originals, [1] kept clones, [2] kept split
 *                    children per draw, [3] split selections, [4] new row count P_new.
 *                    dense_extent = percent_dense * scene_extent, rounded to float; the limit
 *                    big_world_extent = 0.1 * scene_extent applies only when the reference passes a
 *                    max_screen_size (train.py:323), pass a negative value otherwise. split_n is the
 *                    reference's N (only its default 2 is supported, else E3_ERR_UNSUPPORTED).
 *  e3_densify_apply  after the caller has read counts and allocated P_new-row outputs: ONE gather
 *                    launch over all arrays. Output order = the reference's: kept originals, kept
 *                    clones, kept children of the 1st draw, of the 2nd draw, each in source order.
 *                    kind 0: rows copied. kind 1 (xyz, width 3): children get
 *                    xyz + R(rotation) (exp(scaling) * z). kind 2 (scaling, width 3): children get
 *                    log(exp(s) / (0.8 * split_n)). normals = the standard-normal draws z
 *                    [split_n * counts[3], 3] in the reference's order (row c * counts[3] + rank of
 *                    the selected Gaussian), i.e. what torch.normal(mean=0, std=stds) scales
 *                    (gaussian_model.py:358-360). exp_avg / exp_avg_sq (optional): moments follow
 *                    kept originals, new rows get zeros. rotation / scaling = the CURRENT (source)
 *                    [P,4] / [P,3] raw parameters.
 * The caller zeroes xyz_gradient_accum, denom, max_radii2D at the new size (gaussian_model.py:345-347). */
#define E3_DENSIFY_MAX_ARRAYS 8
typedef struct e3_densify_array {
	const float* src;             /* [P, width] */
	float* dst;                   /* [P_new, width] */
	const float* exp_avg_src;     /* optional */
	float* exp_avg_dst;
	const float* exp_avg_sq_src;  /* optional */
	float* exp_avg_sq_dst;
	int width;
	int kind;
} e3_densify_array;
E3_API size_t e3_densify_scratch_bytes(int P);
E3_API int e3_densify_plan(int P, const float* xyz_gradient_accum, const float* denom, const float* scaling,
	const float* opacity, float max_grad, float dense_extent, float min_opacity, float big_world_extent,
	int split_n, void* scratch, size_t scratch_bytes, int* counts, void* stream);
E3_API int e3_densify_apply(int P, int P_new, int num_split_selected, const e3_densify_array* arrays_host,
	int num_arrays, const float* rotation, const float* scaling, const float* normals, int split_n,
	const void* scratch, size_t scratch_bytes, void* stream);

/* GaussianModel.reset_opacity (scene/gaussian_model.py:210-213, train.py:326-327), in place:
 * opacity = inverse_sigmoid(min(sigmoid(opacity), max_opacity)) with max_opacity = 0.01, and the
 * two Adam moments of the opacity group (optional) zeroed as replace_tensor_to_optimizer does. */
E3_API int e3_reset_opacity(int P, float* opacity, float* exp_avg, float* exp_avg_sq, float max_opacity,
	void* stream);

/* dst[i] += scale * src[i] on every rank at once: multicast_dst is the NVLS multicast alias of a
 * symmetric-memory buffer (see e3_event_iteration_args.grad_multicast_delta), src is local. Used for
 * the gradient of the event threshold c, which the loss kernel produces. No counterpart in the
 * (single-GPU) reference. */
E3_API int e3_multimem_add(float* multicast_dst, const float* src, size_t n, float scale, void* stream);

/* Gradient all-reduce of the replicated multi-GPU trainer (no counterpart in the
 * single-GPU reference) as ONE kernel per rank over NVLink 5 / NVSwitch: multicast_ptr is the NVLS
 * multicast alias of a buffer of n floats (n % 4 == 0) that exists in symmetric memory on all `world`
 * ranks. Rank r reduces its 1/world of the buffer inside the switch (multimem.ld_reduce) and writes
 * scale * sum into every rank's copy (multimem.st). The caller issues a cross-rank barrier on the
 * stream before (local gradients complete everywhere) and after (result complete everywhere). */
E3_API int e3_nvls_allreduce(float* multicast_ptr, size_t n, int rank, int world, float scale, void* stream);

/* L1 + SSIM loss of the two non-event training branches, value and backward seed:
 * loss = (1 - lambda_dssim) * L1 + lambda_dssim * (1 - SSIM), with the 11x11 Gaussian-window SSIM
 * of utils/loss_utils.py:359-418 (sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2).
 * gray = 1: the `--gray` branch of train.py:213-223 (l1_loss_gray + ssim_gray on the luma
 * 0.299 R + 0.587 G + 0.114 B, utils/loss_utils.py:18-23,40-48,368-385);
 * gray = 0: the RGB branch of train.py:292-296 (l1_loss + ssim over the 3 channels).
 * image is the render [3,H,W] float; gt the ground-truth frame [3,H,W], float32 or 8-bit
 * (gt_is_u8, /255 applied). scratch: e3_l1_ssim_scratch_bytes() of 8-byte aligned device memory.
 * Outputs: d_image [3,H,W] = upstream * dloss/dimage, out3 = { loss, L1, SSIM }. */
E3_API size_t e3_l1_ssim_scratch_bytes(int width, int height, int gray);
E3_API int e3_l1_ssim_loss(int width, int height, int gray, const float* image, const void* gt,
	int gt_is_u8, float lambda_dssim, float upstream, void* scratch, size_t scratch_bytes,
	float* d_image, float* out3, void* stream);

/* One whole event training iteration in one call: the three renders of train.py:144,159-162, the
 * event loss of :165-203 and loss.backward() (:211) down to the leaf parameters, enqueued without
 * any host synchronisation (e3_rasterize_forward_raw_ws x3, e3_event_loss, e3_rasterize_backward_raw
 * x3). Exists because at event-camera resolutions the ~75 kernels of an iteration are shorter than
 * the host work of seven separate binding calls. Everything is caller-owned and may be reused
 * from call to call.
 * views[0] = intensity view (3 channels), views[1], views[2] = event views `index`, `index+1`
 * (channels 1 = luma, or 3). seed buffers are [channels,H,W]. d_* receive the SUM of the three
 * backward passes (zeroed first, or added to when accumulate != 0; the three views run on three
 * streams and add with vector reductions, so the summation order, like that of the blend's atomics,
 * is not fixed from run to run); d_mean2D (optional) is the screen-space gradient of views[0]. out5 as in e3_event_loss.
 * views[1].channels == 0 selects the shared-first-view form (E3_EVENT_SHARED_FIRST_VIEW): views[1] is not
 * rendered, only its seed buffer ([3,H,W]) is used, views[2].channels must be 1 and d_mean2D is required.
 * After the stream has passed the
 * read-backs, check views[k].num_rendered_host[0] <= binning_capacity (see
 * e3_rasterize_forward_raw_ws). */
typedef struct e3_view {
	const float* viewmatrix;
	const float* projmatrix;
	const float* cam_pos;
	float tan_fovx, tan_fovy;
	int channels;
	char* geometry_buffer;
	char* binning_buffer;
	char* image_buffer;
	float* out_color;
	int* radii;
	float* seed;
	int* num_rendered_host;
} e3_view;
typedef struct e3_event_iteration_args {
	int P, active_sh_degree, width, height;
	unsigned long long binning_capacity;
	float scale_modifier;
	int accumulate;
	const float* background;
	const float *xyz, *features_dc, *features_rest, *opacity_logits, *log_scales, *raw_rotations;
	float *d_xyz, *d_features_dc, *d_features_rest, *d_opacity_logits, *d_log_scales, *d_raw_rotations, *d_mean2D;
	e3_view views[3];
	const void *gt_now, *gt_next, *gt_intensity, *gt_blur;
	int gt_is_u8;
	const float* c;
	float* out5;
	void* loss_scratch;          /* >= 64 bytes, 8-byte aligned */
	void* wait_before_loss_event;/* optional cudaEvent_t the stream waits for right before the loss kernels (e.g. the
	                                upload of the GT frames on another stream, which then overlaps the renders) */
	void* forwards_done_event;   /* optional cudaEvent_t, recorded after the third forward: waiting on it
	                                makes the three num_rendered read-backs visible without draining the stream */
	float loss_upstream;         /* factor on the loss gradient (0 is read as 1): 1 / (view pairs of the optimizer
	                                step) makes the accumulated gradient the one of the MEAN loss */
	long long grad_multicast_delta; /* 0, or the byte offset from each d_* pointer (d_mean2D excepted) to its
	                                NVLS multicast alias: the gradient tensors then live in symmetric memory
	                                mapped on every rank, and the accumulation kernels reduce into ALL ranks'
	                                copies with multimem.red, i.e. the gradient all-reduce of the replicated
	                                multi-GPU trainer (the reference has none) happens inside
	                                them. Requires accumulate != 0: the caller zeroes its copy, passes a
	                                cross-rank barrier, runs the iteration(s), passes a second barrier, and
	                                then holds the sum over all ranks. */
} e3_event_iteration_args;
/* Not re-entrant per device: the three side streams (two views, and one for what needs no render: zeroing the gradient
 * rows, the GT-only part of the loss) and their fork/join events exist once per device, so one
 * host thread per GPU may be inside this call at a time (the one-process-per-GPU model of SURVEY.md 8b). On any
 * return, error returns included, `stream` is ordered behind all work the call gave to its side streams. */
E3_API int e3_event_iteration(const e3_event_iteration_args* args_host, void* stream);

/* Replaces CudaRasterizer::Rasterizer::markVisible (cuda_rasterizer/rasterizer.h:24-29,
 * rasterizer_impl.cu:54-66,141-153): present[i] = view-space z > 0.2. present is uint8[P]. */
E3_API int e3_mark_visible(int P, const float* means3D, const float* viewmatrix,
	const float* projmatrix, uint8_t* present, void* stream);

/* Replaces simple_knn._C.distCUDA2 (call site scene/gaussian_model.py:134; implementation in the
 * un-vendored submodule submodules/simple-knn @ 9302c6d): mean_dist2[i] = mean of the squared
 * Euclidean distances from point i to its 3 nearest OTHER points (exact; coincident points count
 * with distance 0; fewer than 4 points leave FLT_MAX terms in the mean. That last rule is how the
 * upstream kernel is REMEMBERED to behave (its best-distance slots start at FLT_MAX): the source is
 * not in the reference tree and the reference holds no vector for it, so it is not evidence).
 * points is float[P,3]; scratch is caller-owned device memory of e3_dist2_scratch_bytes(P) bytes,
 * 256-byte aligned. */
E3_API size_t e3_dist2_scratch_bytes(int P);
E3_API int e3_dist2(int P, const float* points, float* mean_dist2, void* scratch,
	size_t scratch_bytes, void* stream);

/* Exclusive prefix sum of n uint32 values: out[i] = in[0] + ... + in[i-1]; *total (optional,
 * device) receives the sum of all n. Replaces the cub::DeviceScan::InclusiveSum over
 * tiles_touched and its temp-storage size query (cuda_rasterizer/rasterizer_impl.cu:166,277):
 * the forward uses this kernel for the instance offsets; the reference's inclusive value is
 * out[i] + in[i] and its num_rendered is *total. `scratch` is caller-owned device memory of at
 * least e3_scan_scratch_bytes(n) bytes, 4-byte aligned, contents irrelevant. in == out is allowed. */
E3_API size_t e3_scan_scratch_bytes(size_t n);
E3_API int e3_exclusive_scan(const uint32_t* in, uint32_t* out, size_t n, void* scratch,
	size_t scratch_bytes, uint32_t* total, void* stream);

/* The same prefix sum for callers that cannot supply scratch memory. No counterpart in the
 * reference. Meant for hosts that may not allocate when they call us: a render or serving loop
 * that reserved all device memory up front and bans cudaMalloc in the loop (it synchronises the
 * device), or a plugin host that passes inputs and outputs only. The block sums live in a
 * 32.8 KB array inside the library (present on every GPU that loads it, until exit); the call
 * allocates nothing. n may not exceed E3_NOALLOC_SCAN_MAX, else E3_ERR_TOO_LARGE is returned and
 * nothing runs. The first call on a device resolves the array's device address (do it at
 * start-up, before any stream capture). All calls on a device share the array: calls that
 * overlap on the GPU race with each other, so keep them on one stream or order them with
 * events. Nothing in the array needs to be zero beforehand (every word is written before it is
 * read), so left-over contents of an earlier call are harmless. in == out is allowed. */
#define E3_NOALLOC_SCAN_MAX 16777216
#define E3_ERR_TOO_LARGE (-5)      /* input exceeds a fixed-capacity entry point's maximum */
E3_API size_t e3_exclusive_scan_noalloc_max(void);
E3_API int e3_exclusive_scan_noalloc(const uint32_t* in, uint32_t* out, size_t n,
	uint32_t* total, void* stream);

/* Optional per-stage device timing, used by bench.py for the roofline of the dominant kernel.
 * No counterpart in the reference (it only brackets the whole iteration with two CUDA events,
 * train.py:64-65,95,299). While bit `s` of stage_mask is set, every launch group of stage `s` is
 * bracketed by two cudaEvents recorded on the call's stream; e3_profile_read() waits for the
 * recorded events, adds their elapsed milliseconds to per-stage sums and returns sums and launch
 * counts (arrays of E3_NUM_STAGES). Process-global and meant for one host thread. */
enum {
	E3_STAGE_PREPROCESS_FWD = 0, E3_STAGE_DEPTH_SORT = 1, E3_STAGE_OFFSETS_SCAN = 2, E3_STAGE_EMIT = 3,
	E3_STAGE_TILE_SORT = 4, E3_STAGE_TILE_RANGES = 5, E3_STAGE_RENDER_FWD = 6, E3_STAGE_RENDER_BWD = 7,
	E3_STAGE_PREPROCESS_BWD = 8, E3_STAGE_EVENT_LOSS = 9, E3_STAGE_DIST2 = 10, E3_STAGE_ACTIVATE = 11,
	E3_NUM_STAGES = 16
};
E3_API int e3_profile_enable(unsigned stage_mask);
/* Number of CUDA kernels launched by this library since it was loaded (no reference counterpart). */
E3_API unsigned long long e3_kernel_launches(void);
E3_API int e3_profile_read(double* ms_sum_host, long long* count_host, int reset);

/* Test / inspection hooks: device pointers into the opaque buffers of a finished forward.
 * `which` (geometry): 0 depths f32[P], 1 packed records float4[3P] ({x,y,conic.x,conic.y},
 *   {conic.z,opacity,c0,c1},{c2,-,-,-}), 2 cov3D f32[6P], 3 tiles_touched u32[P],
 *   4 depth_order u32[P], 5 rgb f32[3P], 6 clamped u8[3P].
 * `which` (binning): 0 point_list u32[R] (sorted Gaussian ids = the reference's point_list),
 *   1 sorted tile ids u32[R] (high words of the reference's point_list_keys).
 * `which` (image): 0 final_T f32[N], 1 n_contrib u32[N], 2 ranges uint2[T]. */
E3_API void* e3_geometry_ptr(char* geometry_buffer, int P, int which);
E3_API void* e3_binning_ptr(char* binning_buffer, int num_rendered, int which);
E3_API void* e3_image_ptr(char* image_buffer, int width, int height, int which);

#ifdef __cplusplus
}
#endif
#endif
