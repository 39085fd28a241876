// Binning: instance offsets, (tile | depth) ordering and per-tile ranges. sm_100a.
//
// The reference builds 64-bit keys (tile << 32 | depth bits) for all R Gaussian/tile instances
// and runs one stable radix sort over 32 + log2(tiles) key bits
// (cuda_rasterizer/rasterizer_impl.cu:70-111, 300-308): 6 passes over 12-byte pairs at the
// BASELINE sizes, the largest pure-HBM term of the path. Here the same ORDER is produced with
// far less traffic:
//   1. stable sort of the P Gaussians by their 32 depth-key bits (P << R),
//   2. instances emitted in that order (32-bit tile id + Gaussian id),
//   3. stable sort of the R instances by tile id only (log2(tiles) bits: 2 passes, 8-byte pairs).
// A stable sort by tile of a depth-ordered list equals the stable sort by (tile, depth) with
// ties kept in Gaussian-index order, i.e. exactly what cub::DeviceRadixSort::SortPairs returns
// in the reference (one Gaussian has at most one instance per tile, and equal depths keep index
// order through step 1). The scan and the radix sort are written here (no CUB).
#include "common.cuh"

namespace {

// ---------------- exclusive scan (reduce / scan-of-sums / downsweep) ----------------
__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane)
{
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const uint32_t n = __shfl_up_sync(E3_FULL, v, o);
		if (lane >= o) v += n;
	}
	return v;
}

// exclusive scan of one value per thread across a 256-thread block; returns block total via smem
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp /*[9]*/, uint32_t& total)
{
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t inc = warp_incl_scan(v, lane);
	if (lane == 31) s_warp[w] = inc;
	__syncthreads();
	if (w == 0) {
		uint32_t x = lane < 8 ? s_warp[lane] : 0;
		const uint32_t xi = warp_incl_scan(x, lane);
		if (lane < 8) s_warp[lane] = xi - x;
		if (lane == 7) s_warp[8] = xi;
	}
	__syncthreads();
	total = s_warp[8];
	const uint32_t r = s_warp[w] + inc - v;
	__syncthreads();
	return r;
}

__global__ void __launch_bounds__(E3_SCAN_THREADS) scan_reduce_kernel(const uint32_t* __restrict__ in,
	const uint32_t* __restrict__ idx, size_t n, uint32_t* __restrict__ sums)
{
	__shared__ uint32_t s_warp[9];
	const size_t base = (size_t)blockIdx.x * E3_SCAN_TILE + (size_t)threadIdx.x * E3_SCAN_ITEMS;
	uint32_t v = 0;
#pragma unroll
	for (int i = 0; i < E3_SCAN_ITEMS; i++) {
		const size_t k = base + i;
		if (k < n) v += idx ? in[idx[k]] : in[k];
	}
	uint32_t total;
	block_excl_scan(v, s_warp, total);
	if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: exclusive scan of `sums` in place, grand total to sums[nb] and *total_out
__global__ void __launch_bounds__(E3_SCAN_THREADS) scan_sums_kernel(uint32_t* sums, size_t nb, uint32_t* total_out)
{
	__shared__ uint32_t s_warp[9];
	uint32_t carry = 0;
	for (size_t start = 0; start < nb; start += E3_SCAN_THREADS) {
		const size_t k = start + threadIdx.x;
		const uint32_t v = k < nb ? sums[k] : 0;
		uint32_t total;
		const uint32_t ex = block_excl_scan(v, s_warp, total);
		if (k < nb) sums[k] = carry + ex;
		carry += total;
	}
	if (threadIdx.x == 0) {
		sums[nb] = carry;
		if (total_out) *total_out = carry;
	}
}

__global__ void __launch_bounds__(E3_SCAN_THREADS) scan_down_kernel(const uint32_t* __restrict__ in,
	const uint32_t* __restrict__ idx, size_t n, const uint32_t* __restrict__ sums, uint32_t* __restrict__ out)
{
	__shared__ uint32_t s_warp[9];
	const size_t base = (size_t)blockIdx.x * E3_SCAN_TILE + (size_t)threadIdx.x * E3_SCAN_ITEMS;
	uint32_t x[E3_SCAN_ITEMS];
	uint32_t v = 0;
#pragma unroll
	for (int i = 0; i < E3_SCAN_ITEMS; i++) {
		const size_t k = base + i;
		x[i] = k < n ? (idx ? in[idx[k]] : in[k]) : 0;
		v += x[i];
	}
	uint32_t total;
	uint32_t run = block_excl_scan(v, s_warp, total) + sums[blockIdx.x];
#pragma unroll
	for (int i = 0; i < E3_SCAN_ITEMS; i++) {
		const size_t k = base + i;
		if (k < n) out[k] = run;
		run += x[i];
	}
}

// ---------------- LSD radix sort, one digit per pass ----------------
// counts layout: counts[digit * nblk + block]  (digit-major, so one exclusive scan of the whole
// array yields every block's output base for every digit)
__global__ void __launch_bounds__(E3_SORT_THREADS) radix_hist_kernel(const uint32_t* __restrict__ keys, size_t n,
	int shift, uint32_t mask, uint32_t* __restrict__ counts, uint32_t nblk)
{
	__shared__ uint32_t hist[256];
	hist[threadIdx.x] = 0;
	__syncthreads();
	const size_t start = (size_t)blockIdx.x * E3_SORT_TILE;
	const size_t end = min(n, start + E3_SORT_TILE);
	const int lane = threadIdx.x & 31;
	for (size_t k = start + threadIdx.x; k < start + E3_SORT_TILE; k += E3_SORT_THREADS) {
		const bool valid = k < end;
		const uint32_t d = valid ? ((keys[k] >> shift) & mask) : 0xffffu;
		// warp-aggregate equal digits (neighbouring instances often share a tile)
		const uint32_t peers = __match_any_sync(E3_FULL, d);
		if (valid && lane == (__ffs(peers) - 1)) atomicAdd(&hist[d], __popc(peers));
	}
	__syncthreads();
	if (threadIdx.x <= mask) counts[(size_t)threadIdx.x * nblk + blockIdx.x] = hist[threadIdx.x];
}

__global__ void __launch_bounds__(E3_SORT_THREADS) radix_scatter_kernel(const uint32_t* __restrict__ keys_in,
	const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
	size_t n, int shift, uint32_t mask, const uint32_t* __restrict__ bases, uint32_t nblk)
{
	__shared__ uint32_t wcnt[8][256];
	for (int i = threadIdx.x; i < 8 * 256; i += E3_SORT_THREADS) (&wcnt[0][0])[i] = 0;
	__syncthreads();
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t lt = (1u << lane) - 1u;
	const size_t wstart = (size_t)blockIdx.x * E3_SORT_TILE + (size_t)w * (32 * E3_SORT_ITEMS);
	uint32_t key[E3_SORT_ITEMS];
	uint16_t rank[E3_SORT_ITEMS];
	// items are taken in index order (warp, round, lane), so equal digits keep their order
#pragma unroll
	for (int i = 0; i < E3_SORT_ITEMS; i++) {
		const size_t k = wstart + (size_t)i * 32 + lane;
		const bool valid = k < n;
		key[i] = valid ? keys_in[k] : 0xffffffffu;
		const uint32_t d = valid ? ((key[i] >> shift) & mask) : 0xffffu;
		const uint32_t peers = __match_any_sync(E3_FULL, d);
		const int leader = __ffs(peers) - 1;
		uint32_t base = 0;
		if (valid && lane == leader) { base = wcnt[w][d]; wcnt[w][d] = base + __popc(peers); }
		base = __shfl_sync(E3_FULL, base, leader);
		rank[i] = (uint16_t)(base + __popc(peers & lt));
		__syncwarp();
	}
	__syncthreads();
	// per digit: exclusive scan over the 8 warps, seeded with this block's global base
	if (threadIdx.x <= mask) {
		uint32_t run = bases[(size_t)threadIdx.x * nblk + blockIdx.x];
#pragma unroll
		for (int ww = 0; ww < 8; ww++) { const uint32_t c = wcnt[ww][threadIdx.x]; wcnt[ww][threadIdx.x] = run; run += c; }
	}
	__syncthreads();
#pragma unroll
	for (int i = 0; i < E3_SORT_ITEMS; i++) {
		const size_t k = wstart + (size_t)i * 32 + lane;
		if (k < n) {
			const uint32_t d = (key[i] >> shift) & mask;
			const uint32_t pos = wcnt[w][d] + rank[i];
			keys_out[pos] = key[i];
			vals_out[pos] = vals_in[k];
		}
	}
}

// ---------------- instance emission (duplicateWithKeys, rasterizer_impl.cu:70-111) ----------------
// One thread per Gaussian in DEPTH order. Small footprints are written by the owning lane;
// footprints of more than 32 tiles are written cooperatively by the whole warp, so one large
// splat does not stall 31 idle lanes (the reference's serial inner loop does).
__global__ void __launch_bounds__(256) emit_instances_kernel(int P, const uint32_t* __restrict__ order,
	const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ tiles_touched,
	const uint2* __restrict__ rect, int gx, uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	const int lane = threadIdx.x & 31;
	uint32_t idx = 0, tt = 0, off = 0;
	uint2 rc = make_uint2(0, 0);
	if (k < P) {
		idx = order[k];
		tt = tiles_touched[idx];
		if (tt) { off = offsets[k]; rc = rect[idx]; }
	}
	const uint32_t x0 = rc.x & 0xffff, x1 = rc.x >> 16, y0 = rc.y & 0xffff;
	const uint32_t wdt = x1 - x0;
	if (tt && tt <= 32) {
		uint32_t o = off;
		const uint32_t y1 = rc.y >> 16;
		for (uint32_t y = y0; y < y1; y++)
			for (uint32_t x = x0; x < x1; x++) { tkeys[o] = y * gx + x; vals[o] = idx; o++; }
	}
	uint32_t big = __ballot_sync(E3_FULL, tt > 32);
	while (big) {
		const int src = __ffs(big) - 1;
		big &= big - 1;
		const uint32_t b_idx = __shfl_sync(E3_FULL, idx, src), b_tt = __shfl_sync(E3_FULL, tt, src);
		const uint32_t b_off = __shfl_sync(E3_FULL, off, src), b_x0 = __shfl_sync(E3_FULL, x0, src);
		const uint32_t b_y0 = __shfl_sync(E3_FULL, y0, src), b_w = __shfl_sync(E3_FULL, wdt, src);
		for (uint32_t i = lane; i < b_tt; i += 32) {
			const uint32_t yy = i / b_w, xx = i - yy * b_w;
			tkeys[b_off + i] = (b_y0 + yy) * gx + (b_x0 + xx);
			vals[b_off + i] = b_idx;
		}
	}
}

// identifyTileRanges, rasterizer_impl.cu:116-138, on 32-bit tile ids
__global__ void __launch_bounds__(256) tile_ranges_kernel(size_t L, const uint32_t* __restrict__ tkeys,
	uint2* __restrict__ ranges)
{
	const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= L) return;
	const uint32_t cur = tkeys[idx];
	if (idx == 0) ranges[cur].x = 0;
	else {
		const uint32_t prev = tkeys[idx - 1];
		if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
	}
	if (idx == L - 1) ranges[cur].y = (uint32_t)L;
}

}  // namespace

cudaError_t e3_exclusive_scan_u32(const uint32_t* in, const uint32_t* idx, uint32_t* out, size_t n,
	uint32_t* tmp, uint32_t* total, cudaStream_t st)
{
	if (n == 0) {
		if (total) return cudaMemsetAsync(total, 0, sizeof(uint32_t), st);
		return cudaSuccess;
	}
	const size_t nb = e3_scan_blocks(n);
	scan_reduce_kernel<<<(unsigned)nb, E3_SCAN_THREADS, 0, st>>>(in, idx, n, tmp);
	scan_sums_kernel<<<1, E3_SCAN_THREADS, 0, st>>>(tmp, nb, total);
	scan_down_kernel<<<(unsigned)nb, E3_SCAN_THREADS, 0, st>>>(in, idx, n, tmp, out);
	g_e3_launches += 3;
	return cudaGetLastError();
}

cudaError_t e3_radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], size_t n, int end_bit,
	uint32_t* counts, uint32_t* scan_tmp, int start_buf, int* result_buf, cudaStream_t st)
{
	int cur = start_buf;
	if (n > 0 && end_bit > 0) {
		const int npass = (end_bit + 7) / 8;
		const size_t nblk = e3_sort_blocks(n);
		int lo = 0;
		for (int p = 0; p < npass; p++) {
			// split the bits evenly over the passes (e.g. 12 -> 6 + 6): fewer digits per pass
			// means longer contiguous runs per block in the scatter
			const int nb = (end_bit - lo + (npass - p) - 1) / (npass - p);
			const uint32_t mask = (1u << nb) - 1u;
			const size_t nc = (size_t)(mask + 1) * nblk;
			radix_hist_kernel<<<(unsigned)nblk, E3_SORT_THREADS, 0, st>>>(keys[cur], n, lo, mask, counts, (uint32_t)nblk);
			cudaError_t e = e3_exclusive_scan_u32(counts, nullptr, counts, nc, scan_tmp, nullptr, st);
			if (e != cudaSuccess) return e;
			radix_scatter_kernel<<<(unsigned)nblk, E3_SORT_THREADS, 0, st>>>(keys[cur], vals[cur], keys[cur ^ 1],
				vals[cur ^ 1], n, lo, mask, counts, (uint32_t)nblk);
			g_e3_launches += 2;
			cur ^= 1;
			lo += nb;
		}
	}
	*result_buf = cur;
	return cudaGetLastError();
}

cudaError_t e3_launch_emit_instances(int P, const uint32_t* order, const uint32_t* offsets,
	const uint32_t* tiles_touched, const uint2* rect, int gx, uint32_t* tkeys, uint32_t* vals,
	cudaStream_t st)
{
	emit_instances_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, order, offsets, tiles_touched, rect, gx, tkeys, vals);
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_tile_ranges(size_t R, const uint32_t* tkeys, uint2* ranges, size_t T, cudaStream_t st)
{
	cudaError_t e = cudaMemsetAsync(ranges, 0, T * sizeof(uint2), st);
	if (e != cudaSuccess) return e;
	if (R > 0) { tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, tkeys, ranges); g_e3_launches += 1; }
	return cudaGetLastError();
}
