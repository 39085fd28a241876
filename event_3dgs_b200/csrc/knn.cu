// distCUDA2: mean squared distance of every point to its 3 nearest other points (exact). sm_100a.
//
// Replaces simple_knn._C.distCUDA2 (call site scene/gaussian_model.py:134; the implementation,
// simple-knn @ 9302c6d, is an un-vendored submodule and absent from the reference tree). Its
// contract: best distances start at FLT_MAX, the point itself is excluded by index (coincident
// points count with distance 0), result = (d0 + d1 + d2) / 3 in fp32.
//
// Design: uniform grid over the bulk of the cloud (about 2 points per cell; robust extent, outliers go to the
// border cells), points counting-sorted by
// cell with the library's own radix sort, one thread per point IN CELL ORDER (neighbouring lanes
// walk neighbouring cells, point records are contiguous float4), exact ring expansion: after all
// cells within Chebyshev ring r are scanned the search stops as soon as the third-best distance is
// not larger than the distance to the nearest unscanned cell face. HBM traffic is a few dozen
// bytes per point; the kernel is latency bound on the gather of ~27 cells per query.
#include "common.cuh"
#include <cfloat>

namespace {

struct KnnGrid {
	float lo[3]; float cell[3]; float inv[3]; int g;
	// robust extent (see knn_moments_kernel): count, sum, sum of squares per pass; current trimming box
	double sums[3][7];
	float trim_c[3], trim_w[3];
	unsigned outside[8];       // points outside mean +- kBoxSigmas[i] * sigma
	float box_lo[3], box_hi[3];
};
__constant__ float kBoxSigmas[8] = { 3.f, 4.f, 6.f, 8.f, 12.f, 16.f, 24.f, 32.f };

__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

// bbox[0..2] = min (ordered-uint encoding), bbox[3..5] = max
__global__ void __launch_bounds__(256) knn_bbox_kernel(int P, const float* __restrict__ pts, uint32_t* __restrict__ bbox)
{
	float mn[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, mx[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x)
#pragma unroll
		for (int a = 0; a < 3; a++) { const float v = pts[3 * (size_t)i + a]; mn[a] = fminf(mn[a], v); mx[a] = fmaxf(mx[a], v); }
#pragma unroll
	for (int a = 0; a < 3; a++) {
		for (int o = 16; o; o >>= 1) {
			mn[a] = fminf(mn[a], __shfl_xor_sync(E3_FULL, mn[a], o));
			mx[a] = fmaxf(mx[a], __shfl_xor_sync(E3_FULL, mx[a], o));
		}
		if ((threadIdx.x & 31) == 0) { atomicMin(bbox + a, f2ord(mn[a])); atomicMax(bbox + 3 + a, f2ord(mx[a])); }
	}
}

// The grid must follow the BULK of the cloud, not its bounding box: COLMAP clouds carry far outliers, and a box
// stretched by them puts almost every point into a few cells (each query then scans its cell serially: quadratic).
// Three passes of mean / standard deviation, the second and third over the points inside mean +- 2.5 sigma of the
// pass before (trimmed moments; 1 % of outliers at 100x the extent are gone after one trimming). The grid then spans
// mean +- k sigma (k chosen by knn_grid_kernel), clipped to the bounding box; points outside fall into the border cells
// (cell_coord clamps), which the exact search treats like any other cell, so the result does not depend on the box --
// only the running time.
__global__ void __launch_bounds__(256) knn_moments_kernel(int P, const float* __restrict__ pts, KnnGrid* __restrict__ grid, int pass)
{
	const bool trim = pass > 0;
	float c[3] = { 0.f, 0.f, 0.f }, w[3] = { 0.f, 0.f, 0.f };
	if (trim) for (int a = 0; a < 3; a++) { c[a] = grid->trim_c[a]; w[a] = grid->trim_w[a]; }
	double acc[7] = { 0, 0, 0, 0, 0, 0, 0 };
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
		const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
		if (trim && !(fabsf(x - c[0]) <= w[0] && fabsf(y - c[1]) <= w[1] && fabsf(z - c[2]) <= w[2])) continue;
		acc[0] += 1.0; acc[1] += x; acc[2] += y; acc[3] += z;
		acc[4] += (double)x * x; acc[5] += (double)y * y; acc[6] += (double)z * z;
	}
#pragma unroll
	for (int k = 0; k < 7; k++) {
		for (int o = 16; o; o >>= 1) acc[k] += __shfl_xor_sync(E3_FULL, acc[k], o);
		if ((threadIdx.x & 31) == 0 && acc[k] != 0.0) atomicAdd(&grid->sums[pass][k], acc[k]);
	}
}

__global__ void knn_trim_kernel(KnnGrid* __restrict__ grid, int pass)
{
	if (threadIdx.x != 0 || blockIdx.x != 0) return;
	const double n = grid->sums[pass][0] > 0.0 ? grid->sums[pass][0] : 1.0;
	for (int a = 0; a < 3; a++) {
		const double mean = grid->sums[pass][1 + a] / n;
		const double var = grid->sums[pass][4 + a] / n - mean * mean;
		grid->trim_c[a] = (float)mean;
		grid->trim_w[a] = (float)(sqrt(var > 0.0 ? var : 0.0) * (pass < 2 ? 2.5 : 1.0));      // after the last pass: sigma itself
	}
}

// how many points lie outside mean +- k sigma, for the candidate k of kBoxSigmas
__global__ void __launch_bounds__(256) knn_count_outside_kernel(int P, const float* __restrict__ pts, KnnGrid* __restrict__ grid)
{
	float c[3], w[3];
	for (int a = 0; a < 3; a++) { c[a] = grid->trim_c[a]; w[a] = grid->trim_w[a]; }
	unsigned cnt[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
		const float dx = fabsf(pts[3 * (size_t)i] - c[0]), dy = fabsf(pts[3 * (size_t)i + 1] - c[1]), dz = fabsf(pts[3 * (size_t)i + 2] - c[2]);
#pragma unroll
		for (int k = 0; k < 8; k++) cnt[k] += !(dx <= kBoxSigmas[k] * w[0] && dy <= kBoxSigmas[k] * w[1] && dz <= kBoxSigmas[k] * w[2]);
	}
#pragma unroll
	for (int k = 0; k < 8; k++) {
		for (int o = 16; o; o >>= 1) cnt[k] += __shfl_xor_sync(E3_FULL, cnt[k], o);
		if ((threadIdx.x & 31) == 0 && cnt[k]) atomicAdd(&grid->outside[k], cnt[k]);
	}
}

__global__ void knn_grid_kernel(const uint32_t* __restrict__ bbox, int g, int P, KnnGrid* __restrict__ grid)
{
	if (threadIdx.x != 0 || blockIdx.x != 0) return;
	grid->g = g;
	// the tightest box mean +- k sigma (of the trimmed cloud) that leaves at most 1 % of the points outside: those are
	// answered by knn_outside_kernel, whose cost is (points outside) x P
	int pick = 7;
	for (int k = 7; k >= 0; k--) if (grid->outside[k] * 100ull <= (unsigned long long)P) pick = k;
	for (int a = 0; a < 3; a++) {
		const float half = kBoxSigmas[pick] * grid->trim_w[a];
		const float lo = fmaxf(ord2f(bbox[a]), grid->trim_c[a] - half), hi = fminf(ord2f(bbox[3 + a]), grid->trim_c[a] + half);
		grid->box_lo[a] = grid->trim_c[a] - half; grid->box_hi[a] = grid->trim_c[a] + half;
		const float ext = hi - lo;
		grid->lo[a] = lo;
		grid->cell[a] = ext > 0.f ? ext / g : 0.f;
		grid->inv[a] = ext > 0.f ? g / ext : 0.f;
	}
}

__device__ __forceinline__ int cell_coord(float v, float lo, float inv, int g)
{
	return min(g - 1, max(0, (int)((v - lo) * inv)));
}

__global__ void __launch_bounds__(256) knn_cell_kernel(int P, const float* __restrict__ pts, const KnnGrid* __restrict__ grid,
	uint32_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P) return;
	const KnnGrid G = *grid;
	const int ix = cell_coord(pts[3 * (size_t)i], G.lo[0], G.inv[0], G.g);
	const int iy = cell_coord(pts[3 * (size_t)i + 1], G.lo[1], G.inv[1], G.g);
	const int iz = cell_coord(pts[3 * (size_t)i + 2], G.lo[2], G.inv[2], G.g);
	keys[i] = (uint32_t)((iz * G.g + iy) * G.g + ix);
	vals[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(256) knn_gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
	float4* __restrict__ sorted)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= P) return;
	const uint32_t i = order[k];
	sorted[k] = make_float4(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2], __uint_as_float(i));
}

// outside the box the grid was laid over (same comparison as knn_count_outside_kernel's)
__device__ __forceinline__ bool knn_outside(const KnnGrid& G, float x, float y, float z)
{
	return !(x >= G.box_lo[0] && x <= G.box_hi[0] && y >= G.box_lo[1] && y <= G.box_hi[1] && z >= G.box_lo[2] && z <= G.box_hi[2]);
}

__device__ __forceinline__ void update3(float d, float& b0, float& b1, float& b2)
{
	if (d < b2) {
		if (d < b1) { b2 = b1; if (d < b0) { b1 = b0; b0 = d; } else b1 = d; }
		else b2 = d;
	}
}

__global__ void __launch_bounds__(128) knn_query_kernel(int P, const float4* __restrict__ sorted, const uint2* __restrict__ ranges,
	const KnnGrid* __restrict__ grid, float* __restrict__ out)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= P) return;
	const KnnGrid G = *grid;
	const int g = G.g;
	const float4 q = sorted[k];
	if (knn_outside(G, q.x, q.y, q.z)) return;        // answered by knn_outside_kernel
	const int c[3] = { cell_coord(q.x, G.lo[0], G.inv[0], g), cell_coord(q.y, G.lo[1], G.inv[1], g), cell_coord(q.z, G.lo[2], G.inv[2], g) };
	const float qa[3] = { q.x, q.y, q.z };
	// an axis with zero extent has all its points in cell 0: treat it as one cell wide
	const int gd[3] = { G.cell[0] > 0.f ? g : 1, G.cell[1] > 0.f ? g : 1, G.cell[2] > 0.f ? g : 1 };
	float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
	for (int r = 1; r <= g; r++) {
		const int x0 = max(0, c[0] - r), x1 = min(gd[0] - 1, c[0] + r);
		const int y0 = max(0, c[1] - r), y1 = min(gd[1] - 1, c[1] + r);
		const int z0 = max(0, c[2] - r), z1 = min(gd[2] - 1, c[2] + r);
		for (int z = z0; z <= z1; z++)
			for (int y = y0; y <= y1; y++) {
				const bool shell_zy = (r == 1) || abs(z - c[2]) == r || abs(y - c[1]) == r;
				for (int x = x0; x <= x1; x++) {
					// ring r > 1 adds only the shell of the cube; its interior was scanned before
					if (!shell_zy && abs(x - c[0]) != r) continue;
					const uint2 rg = __ldg(ranges + ((size_t)z * g + y) * g + x);
					for (uint32_t j = rg.x; j < rg.y; j++) {
						if ((int)j == k) continue;
						const float4 p = __ldg(sorted + j);
						const float dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
						update3(dx * dx + dy * dy + dz * dz, b0, b1, b2);
					}
				}
			}
		// distance from q to the nearest face of the scanned cube that still has cells behind it
		float dmin = FLT_MAX;
		bool open = false;
#pragma unroll
		for (int a = 0; a < 3; a++) {
			if (c[a] - r > 0) { open = true; dmin = fminf(dmin, qa[a] - (G.lo[a] + (c[a] - r) * G.cell[a])); }
			if (c[a] + r < gd[a] - 1) { open = true; dmin = fminf(dmin, (G.lo[a] + (c[a] + r + 1) * G.cell[a]) - qa[a]); }
		}
		if (!open) break;
		// conservative against rounding of the face positions
		dmin = fmaxf(0.f, dmin) * 0.9999f;
		if (b2 <= dmin * dmin) break;
	}
	out[__float_as_uint(q.w)] = (b0 + b1 + b2) / 3.0f;
}

// The (at most 1 % of the) points outside the grid's box: one warp per point scans ALL points. A ring search from a
// border cell would walk through the whole grid for them (their neighbours are other far points, many cell widths
// away). Same distance expression and the same three-best rule as the grid search, so the result is the exact 3-NN.
__global__ void __launch_bounds__(256) knn_outside_kernel(int P, const float4* __restrict__ sorted, const KnnGrid* __restrict__ grid,
	float* __restrict__ out)
{
	const int k = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
	if (k >= P) return;
	const float4 q = sorted[k];
	if (!knn_outside(*grid, q.x, q.y, q.z)) return;
	float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
	for (int j = lane; j < P; j += 32) {
		if (j == k) continue;
		const float4 p = __ldg(sorted + j);
		const float dx = p.x - q.x, dy = p.y - q.y, dz = p.z - q.z;
		update3(dx * dx + dy * dy + dz * dz, b0, b1, b2);
	}
	// the three smallest over the 32 lanes' sorted triples
	float r[3];
#pragma unroll
	for (int t = 0; t < 3; t++) {
		float m = b0;
		for (int o = 16; o; o >>= 1) m = fminf(m, __shfl_xor_sync(E3_FULL, m, o));
		const int owner = __ffs(__ballot_sync(E3_FULL, b0 == m)) - 1;
		if (lane == owner) { b0 = b1; b1 = b2; b2 = FLT_MAX; }
		r[t] = m;
	}
	if (lane == 0) out[__float_as_uint(q.w)] = (r[0] + r[1] + r[2]) / 3.0f;
}

}  // namespace

int e3_knn_grid_dim(int P)
{
	int g = (int)ceil(cbrt((double)P / 2.0));
	return g < 1 ? 1 : (g > 128 ? 128 : g);
}

size_t e3_knn_scratch_layout(int P, uint32_t** keys, uint32_t** vals, uint32_t** counts, uint32_t** scan_tmp, uint2** ranges,
	float4** sorted, uint32_t** bbox, void** grid, char* base)
{
	const int g = e3_knn_grid_dim(P);
	const size_t ncell = (size_t)g * g * g;
	char* p = base;
	uint32_t *k0, *k1, *v0, *v1, *cn, *stp, *bb; uint2* rg; float4* so; KnnGrid* gr;
	e3_carve(p, k0, (size_t)P); e3_carve(p, k1, (size_t)P); e3_carve(p, v0, (size_t)P); e3_carve(p, v1, (size_t)P);
	e3_carve(p, cn, e3_sort_scratch_words((size_t)P)); e3_carve(p, stp, 8);
	e3_carve(p, rg, ncell); e3_carve(p, so, (size_t)P); e3_carve(p, bb, 8); e3_carve(p, gr, 1);
	if (keys) { keys[0] = k0; keys[1] = k1; vals[0] = v0; vals[1] = v1; *counts = cn; *scan_tmp = stp; *ranges = rg; *sorted = so; *bbox = bb; *grid = gr; }
	return (size_t)(p - base) + 256;
}

cudaError_t e3_launch_dist2(int P, const float* pts, float* out, char* scratch, cudaStream_t st)
{
	uint32_t *keys[2], *vals[2], *counts, *scan_tmp, *bbox; uint2* ranges; float4* sorted; void* gridv;
	e3_knn_scratch_layout(P, keys, vals, &counts, &scan_tmp, &ranges, &sorted, &bbox, &gridv, scratch);
	KnnGrid* grid = static_cast<KnnGrid*>(gridv);
	const int g = e3_knn_grid_dim(P);
	const size_t ncell = (size_t)g * g * g;
	int bits = 0;
	while (((size_t)1 << bits) < ncell) bits++;
	cudaError_t e;
	if ((e = cudaMemsetAsync(bbox, 0xff, 3 * sizeof(uint32_t), st)) != cudaSuccess) return e;
	if ((e = cudaMemsetAsync(bbox + 3, 0, 3 * sizeof(uint32_t), st)) != cudaSuccess) return e;
	const int nb = (P + 255) / 256;
	if ((e = cudaMemsetAsync(grid, 0, sizeof(KnnGrid), st)) != cudaSuccess) return e;
	knn_bbox_kernel<<<nb < 592 ? nb : 592, 256, 0, st>>>(P, pts, bbox);
	for (int pass = 0; pass < 3; pass++) {
		knn_moments_kernel<<<nb < 592 ? nb : 592, 256, 0, st>>>(P, pts, grid, pass);
		knn_trim_kernel<<<1, 32, 0, st>>>(grid, pass);
	}
	knn_count_outside_kernel<<<nb < 592 ? nb : 592, 256, 0, st>>>(P, pts, grid);
	knn_grid_kernel<<<1, 32, 0, st>>>(bbox, g, P, grid);
	knn_cell_kernel<<<nb, 256, 0, st>>>(P, pts, grid, keys[0], vals[0]);
	g_e3_launches += 10;
	int res = 0;
	if ((e = e3_radix_sort_pairs(keys, vals, (size_t)P, bits, counts, scan_tmp, 0, &res, st)) != cudaSuccess) return e;
	if ((e = e3_launch_tile_ranges((size_t)P, keys[res], ranges, ncell, st)) != cudaSuccess) return e;
	knn_gather_kernel<<<nb, 256, 0, st>>>(P, pts, vals[res], sorted);
	knn_query_kernel<<<(P + 127) / 128, 128, 0, st>>>(P, sorted, ranges, grid, out);
	knn_outside_kernel<<<(unsigned)(((size_t)P * 32 + 255) / 256), 256, 0, st>>>(P, sorted, grid, out);
	g_e3_launches += 3;
	return cudaGetLastError();
}
