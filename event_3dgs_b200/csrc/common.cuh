// Shared device helpers and private state layouts of libevent3dgs (sm_100a).
#pragma once
#include <cstdint>
#include <cstddef>
#include <atomic>
#include <cuda_runtime.h>

#define E3_TILE_X 16
#define E3_TILE_Y 16
#define E3_TILE_PIX 256
#define E3_FULL 0xffffffffu

// Per-Gaussian gradient accumulator written by the backward blend and consumed by the
// backward preprocess: 12 floats (48 B, 16-byte aligned) so that one warp's reduced values
// land in one or two 32-byte sectors.
// The blend leaves sums that lack their per-Gaussian factors (g = G dL/dalpha, d = centre - pixel):
//   0 sum g   1,2 -sum g dpower/ddx, -g dpower/ddy   3,4,5 sum g dx^2, g dx dy, g dy^2   6,7,8 dL/dcolour
//   9,10 -sum g dpower/ddx, -g dpower/ddy of the second seed (dual backward blend)
// and the backward preprocess finishes them (preprocess.cu).
#define E3_ACC_STRIDE 12

static inline size_t e3_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <typename T>
static inline void e3_carve(char*& p, T*& out, size_t count)
{
	p = reinterpret_cast<char*>(e3_align_up(reinterpret_cast<size_t>(p), 256));
	out = reinterpret_cast<T*>(p);
	p += count * sizeof(T);
}

// Radix-sort scratch: `counts` holds one digit histogram per sort block.
#define E3_SORT_THREADS 256
#ifndef E3_SORT_ITEMS
#define E3_SORT_ITEMS 16
#endif
#define E3_SORT_TILE (E3_SORT_THREADS * E3_SORT_ITEMS)
#define E3_SCAN_THREADS 256
#define E3_SCAN_ITEMS 8
#define E3_SCAN_TILE (E3_SCAN_THREADS * E3_SCAN_ITEMS)

static inline size_t e3_sort_blocks(size_t n) { return (n + E3_SORT_TILE - 1) / E3_SORT_TILE; }
// Look-back descriptors are two-level: one per tile and one per group of E3_SORT_GROUP tiles.
#define E3_SORT_GROUP 32
static inline size_t e3_sort_groups(size_t n) { return (e3_sort_blocks(n) + E3_SORT_GROUP - 1) / E3_SORT_GROUP; }
static inline size_t e3_sort_desc_words(size_t n) { return (e3_sort_blocks(n) + e3_sort_groups(n)) * 256; }
// uint32 words of radix-sort scratch: 4 global histograms, tile counters, look-back descriptors of 4 passes
static inline size_t e3_sort_scratch_words(size_t n) { return 4 * 256 + 8 + 4 * e3_sort_desc_words(n); }
static inline size_t e3_scan_blocks(size_t n) { return (n + E3_SCAN_TILE - 1) / E3_SCAN_TILE; }

struct E3GeomState {
	float* depths;            // [P] view-space z
	float4* rec;              // [3P] packed per-Gaussian record
	float* cov3D;             // [6P]
	float* rgb;               // [3P] SH colours (only when shs given)
	uint8_t* clamped;         // [3P]
	uint32_t* tiles_touched;  // [P]
	uint2* rect;              // [P] (xmin | xmax<<16, ymin | ymax<<16)
	uint32_t* dkey[2];        // [P] depth keys, ping-pong
	uint32_t* dval[2];        // [P] Gaussian ids in depth order, ping-pong
	uint32_t* offsets;        // [P] exclusive instance offsets in depth order
	uint32_t* sort_counts;    // [e3_sort_scratch_words(P)]
	uint32_t* scan_tmp;       // [scan_blocks(P) + 2]
	float* acc;               // [P * E3_ACC_STRIDE] backward accumulators
	int* flags;               // [4]: 0 num_rendered, 1 prefiltered-violation, 2 (as float) luma of the background
	static E3GeomState carve(char* p, size_t P, char** end = nullptr)
	{
		E3GeomState s;
		e3_carve(p, s.depths, P);
		e3_carve(p, s.rec, 3 * P);
		e3_carve(p, s.cov3D, 6 * P);
		e3_carve(p, s.rgb, 3 * P);
		e3_carve(p, s.clamped, 3 * P);
		e3_carve(p, s.tiles_touched, P);
		e3_carve(p, s.rect, P);
		e3_carve(p, s.dkey[0], P);
		e3_carve(p, s.dkey[1], P);
		e3_carve(p, s.dval[0], P);
		e3_carve(p, s.dval[1], P);
		e3_carve(p, s.offsets, P);
		e3_carve(p, s.sort_counts, e3_sort_scratch_words(P));
		e3_carve(p, s.scan_tmp, e3_scan_blocks(P) + 2);
		e3_carve(p, s.acc, P * E3_ACC_STRIDE);
		e3_carve(p, s.flags, 4);
		if (end) *end = p;
		return s;
	}
};

struct E3ImageState {
	float* final_T;       // [N]
	uint32_t* n_contrib;  // [N]
	uint2* ranges;        // [T]
	uint32_t* tile_order; // [T] tile ids, longest list first
	static E3ImageState carve(char* p, size_t N, size_t T, char** end = nullptr)
	{
		E3ImageState s;
		e3_carve(p, s.final_T, N);
		e3_carve(p, s.n_contrib, N);
		e3_carve(p, s.ranges, T);
		e3_carve(p, s.tile_order, T);
		if (end) *end = p;
		return s;
	}
};

struct E3BinState {
	uint32_t* tkey[2];      // [R] tile ids, ping-pong
	uint32_t* val[2];       // [R] Gaussian ids, ping-pong
	uint32_t* sort_counts;  // [e3_sort_scratch_words(R)]
	uint32_t* scan_tmp;     // unused
	static E3BinState carve(char* p, size_t R, char** end = nullptr)
	{
		E3BinState s;
		e3_carve(p, s.tkey[0], R);
		e3_carve(p, s.tkey[1], R);
		e3_carve(p, s.val[0], R);
		e3_carve(p, s.val[1], R);
		e3_carve(p, s.sort_counts, e3_sort_scratch_words(R));
		e3_carve(p, s.scan_tmp, 8);
		if (end) *end = p;
		return s;
	}
};

// number of kernels this library has launched (bench.py reports it as gpu_launches)
extern std::atomic<unsigned long long> g_e3_launches;

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device setting: done once per (kernel, device), with
// the launching device current (a process may drive several GPUs from several threads).
cudaError_t e3_allow_dynamic_smem_impl(const void* kernel, size_t bytes);      // api.cu
template <typename K>
static inline cudaError_t e3_allow_dynamic_smem(K kernel, size_t bytes)
{
	return e3_allow_dynamic_smem_impl(reinterpret_cast<const void*>(kernel), bytes);
}

// ---- internal launchers (defined in the .cu files) ----
struct E3Frame {
	int P, D, M, C, W, H, gx, gy;
	float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
	// != 0: byte offset from a gradient pointer to its NVLS multicast alias; accumulating gradient
	// writes then go out as multimem.red (reduced into every rank's copy by the NVSwitch)
	long long mc_delta = 0;
	// the screen-space gradient handed out (d_mean2D) comes from accumulator slots 9, 10 (the second seed
	// of a dual backward blend) instead of slots 0, 1
	bool mean2d_from_second_seed = false;
};

cudaError_t e3_launch_preprocess(const E3Frame& f, const float* means3D, const float* scales,
	const float* rotations, const float* opacities, const float* shs, const float* f_rest, bool raw,
	const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix, const float* projmatrix,
	const float* cam_pos, bool prefiltered, int* radii, E3GeomState& g, const float* bg3, cudaStream_t st);
cudaError_t e3_launch_mark_visible(int P, const float* means3D, const float* viewmatrix,
	uint8_t* present, cudaStream_t st);
cudaError_t e3_launch_preprocess_backward(const E3Frame& f, const float* means3D,
	const int* radii, const float* shs, const float* f_rest, const float* opacities_raw, bool raw, bool accumulate,
	const float* scales, const float* rotations,
	const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
	const float* campos, const E3GeomState& g, float* dL_dmean2D, float* dL_dconic,
	float* dL_dopacity, float* dL_dcolor, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* d_frest,
	float* dL_dscale, float* dL_drot, cudaStream_t st);
cudaError_t e3_launch_event_gt_fraction(size_t N, const void* gt_now, const void* gt_next, int gt_is_u8, double* sums, cudaStream_t st);
cudaError_t e3_launch_event_loss(size_t N, const float* image, const float* image_now, const float* image_next,
	const void* gt_now, const void* gt_next, const void* gt_int, const void* gt_blur, int gt_is_u8, const float* c_ptr,
	float upstream, int event_channels, double* sums, float* d_image, float* d_now, float* d_next, float* out, cudaStream_t st);

// out[i] = sum_{j<i} in[idx ? idx[j] : j]; total (optional, device) = sum of all
cudaError_t e3_exclusive_scan_u32(const uint32_t* in, const uint32_t* idx, uint32_t* out,
	size_t n, uint32_t* tmp, uint32_t* total, cudaStream_t st);
// Stable LSD radix sort of (key,val) pairs over key bits [0,end_bit). The input is read from
// ping-pong buffer `start_buf`; *result_buf receives the index (0/1) of the buffer that holds
// the result. e3_radix_passes() tells how many passes (buffer flips) that takes.
// all_ones_is_padding (32-bit keys): entries with key 0xffffffff are padding whose position in the result is
// irrelevant to the caller; they end up behind all other keys, but not necessarily in index order among keys
// that share their low bits.
static inline int e3_radix_passes(int end_bit) { return (end_bit + 7) / 8; }
cudaError_t e3_radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], size_t n, int end_bit,
	uint32_t* counts, uint32_t* scan_tmp, int start_buf, int* result_buf, cudaStream_t st, const uint32_t* n_dev = nullptr,
	bool all_ones_is_padding = false);
cudaError_t e3_launch_emit_instances(int P, const uint32_t* order, const uint32_t* offsets,
	const uint32_t* tiles_touched, const uint2* rect, int gx, uint32_t* tkeys, uint32_t* vals, uint32_t cap,
	cudaStream_t st);
cudaError_t e3_launch_tile_ranges(size_t R, const uint32_t* tkeys, uint2* ranges, size_t T,
	cudaStream_t st, const uint32_t* n_dev = nullptr);

cudaError_t e3_launch_tile_order(const uint2* ranges, int T, uint32_t* order, cudaStream_t st);
cudaError_t e3_launch_render_forward(const E3Frame& f, const uint2* ranges, const uint32_t* tile_order,
	const uint32_t* point_list, const float4* rec, const float* bg, float* final_T,
	uint32_t* n_contrib, float* out_color, cudaStream_t st, bool exact = true);
cudaError_t e3_launch_render_backward(const E3Frame& f, const uint2* ranges, const uint32_t* tile_order,
	const uint32_t* point_list, const float4* rec, const float* bg, const float* final_T,
	const uint32_t* n_contrib, const float* dL_dpix, float* acc, cudaStream_t st, const float* dL_dpix2 = nullptr,
	bool exact = false);

// knn.cu
int e3_knn_grid_dim(int P);
size_t e3_knn_scratch_layout(int P, uint32_t** keys, uint32_t** vals, uint32_t** counts, uint32_t** scan_tmp, uint2** ranges,
	float4** sorted, uint32_t** bbox, void** grid, char* base);
cudaError_t e3_launch_dist2(int P, const float* pts, float* out, char* scratch, cudaStream_t st);

// ssim.cu
size_t e3_l1_ssim_scratch_words(int W, int H, bool gray);
cudaError_t e3_launch_l1_ssim(int W, int H, bool gray, const float* img, const void* gt, bool gt_is_u8, float lambda,
	float upstream, float* scratch, float* d_image, float* out3, cudaStream_t st);
