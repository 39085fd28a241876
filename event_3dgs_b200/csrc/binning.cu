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

// ---------------- LSD radix sort, "onesweep" style ----------------
// One histogram kernel for all passes, then ONE kernel per pass: every block ranks its tile
// locally, obtains the number of equal digits in all earlier tiles by decoupled look-back
// (single-word flag|count descriptors), reorders the tile in shared memory and writes each digit
// run to global memory with consecutive threads on consecutive addresses. Stable.
// Scratch (uint32 words): ghist[4][256] | tile_counter[8] | per pass: tile_desc[ntiles][256], group_desc[ngroups][256].
#ifndef E3_SORT_LOOKBACK
#define E3_SORT_LOOKBACK 16
#endif
#define E3_LB_AGG 0x40000000u
#define E3_LB_PREFIX 0x80000000u
#define E3_LB_MASK 0x3fffffffu

struct RadixPlan { int npass; int shift[4]; int bits[4]; };

// n_dev (optional): the element count lives on the device (no host round trip); `n` is then the
// capacity the launch was sized for and the effective count is min(*n_dev, n).
__global__ void __launch_bounds__(256) radix_hist_all_kernel(const uint32_t* __restrict__ keys, size_t n, RadixPlan plan,
	uint32_t* __restrict__ ghist, const uint32_t* __restrict__ n_dev)
{
	if (n_dev) n = min(n, (size_t)*n_dev);
	__shared__ uint32_t hist[4][256];
	for (int i = threadIdx.x; i < 4 * 256; i += 256) (&hist[0][0])[i] = 0;
	__syncthreads();
	for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (size_t)gridDim.x * 256) {
		const uint32_t key = keys[k];
#pragma unroll
		for (int p = 0; p < 4; p++)
			if (p < plan.npass) atomicAdd(&hist[p][(key >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u)], 1u);
	}
	__syncthreads();
	for (int p = 0; p < plan.npass; p++) {
		const uint32_t c = hist[p][threadIdx.x];
		if (c) atomicAdd(&ghist[p * 256 + threadIdx.x], c);
	}
}

#ifndef E3_SORT_MIN_BLOCKS
#define E3_SORT_MIN_BLOCKS 2
#endif
__global__ void __launch_bounds__(E3_SORT_THREADS, E3_SORT_MIN_BLOCKS) radix_onesweep_kernel(const uint32_t* __restrict__ keys_in,
	const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
	size_t n, int shift, uint32_t mask, const uint32_t* __restrict__ ghist, volatile uint32_t* lookback,
	uint32_t* __restrict__ tile_counter, const uint32_t* __restrict__ n_dev, int padding_bin)
{
	const uint32_t ntiles = gridDim.x;     // the launch is sized for the capacity, the descriptor arrays with it
	if (n_dev) n = min(n, (size_t)*n_dev);
	// All keys agree on this digit (the global histogram has a single non-empty bin): the stable pass
	// is the identity permutation, so the tile is copied to the other buffer without ranking or
	// look-back. Typical for the top byte of the depth keys (sign + 7 exponent bits: all depths of a
	// scene inside [2^k, 4 * 2^k)).
	// padding_bin >= 0 (last pass only): the keys of that bin are all-ones padding whose place in the result does not
	// matter (the depth keys of culled Gaussians). The lower passes have already put them behind everything else, so
	// the pass is the identity as well when the real keys agree on the digit.
	if (__syncthreads_or(threadIdx.x <= mask && n > 0 && (ghist[threadIdx.x] == (uint32_t)n ||
			(padding_bin >= 0 && ghist[threadIdx.x] + ghist[padding_bin] == (uint32_t)n)))) {
		const size_t t0 = (size_t)blockIdx.x * E3_SORT_TILE;
		for (size_t k = t0 + threadIdx.x; k < min(n, t0 + (size_t)E3_SORT_TILE); k += E3_SORT_THREADS) {
			keys_out[k] = keys_in[k];
			vals_out[k] = vals_in[k];
		}
		return;
	}
	__shared__ uint32_t wcnt[8][256];
	__shared__ uint32_t s_keys[E3_SORT_TILE];
	__shared__ uint32_t s_vals[E3_SORT_TILE];
	__shared__ uint32_t s_gpos[256];      // global position of the first element of each digit run of this tile
	__shared__ uint32_t s_lstart[256];    // position of that run inside the tile
	__shared__ uint32_t s_warp[9];
	__shared__ uint32_t s_tile;
	if (threadIdx.x == 0) s_tile = atomicAdd(tile_counter, 1u);
	for (int i = threadIdx.x; i < 8 * 256; i += E3_SORT_THREADS) (&wcnt[0][0])[i] = 0;
	__syncthreads();
	const uint32_t tile = s_tile;
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t lt = (1u << lane) - 1u;
	const size_t tstart = (size_t)tile * E3_SORT_TILE;
	if (tstart >= n) return;       // launch sized for a capacity: tiles past the data have nothing to do (block-uniform)
	const size_t wstart = tstart + (size_t)w * (32 * E3_SORT_ITEMS);
	const uint32_t tcount = (uint32_t)min((size_t)E3_SORT_TILE, n - tstart);
	uint32_t key[E3_SORT_ITEMS], val[E3_SORT_ITEMS];
	uint16_t rank[E3_SORT_ITEMS];
#pragma unroll
	for (int i = 0; i < E3_SORT_ITEMS; i++) {
		const size_t k = wstart + (size_t)i * 32 + lane;
		const bool valid = k < n;
		key[i] = valid ? keys_in[k] : 0xffffffffu;
		val[i] = valid ? vals_in[k] : 0u;
	}
	// items are ranked in index order (warp, round, lane), so equal digits keep their order
#pragma unroll
	for (int i = 0; i < E3_SORT_ITEMS; i++) {
		const bool valid = wstart + (size_t)i * 32 + lane < n;
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
	// per digit (thread d): tile total, look-back over earlier tiles, warp offsets
	uint32_t total = 0;
	const bool mine = threadIdx.x <= mask;
	volatile uint32_t* tile_desc = lookback;                                   // [ntiles][256], aggregate only
	volatile uint32_t* group_desc = lookback + (size_t)ntiles * 256;           // [ngroups][256], aggregate or inclusive prefix
	const uint32_t grp = tile / E3_SORT_GROUP, rin = tile % E3_SORT_GROUP;
	if (mine) {
#pragma unroll
		for (int ww = 0; ww < 8; ww++) { const uint32_t c = wcnt[ww][threadIdx.x]; wcnt[ww][threadIdx.x] = total; total += c; }
		tile_desc[(size_t)tile * 256 + threadIdx.x] = total | E3_LB_AGG;
	}
	// exclusive scans over the digits: global bases from the histogram, local run starts from the totals
	uint32_t tot_all;
	const uint32_t gb = block_excl_scan(mine ? ghist[threadIdx.x] : 0u, s_warp, tot_all);
	const uint32_t ls = block_excl_scan(total, s_warp, tot_all);
	if (mine) {
		// Two-level look-back. All tiles of a launch are co-resident (<= 2 per SM), so a flat look-back
		// would walk over almost every earlier tile (O(tiles^2) descriptor reads). Instead: sum the
		// aggregates of the earlier tiles of the own group (published before any look-back starts, so
		// there is nothing to wait for), then walk the GROUP descriptors, which the last tile of each
		// group publishes first as aggregate and then as inclusive prefix.
		uint32_t in_group = 0;
		constexpr int LB = E3_SORT_LOOKBACK;
		for (int k0 = 0; k0 < (int)rin; k0 += LB) {
			uint32_t v[LB];
#pragma unroll
			for (int k = 0; k < LB; k++) v[k] = (k0 + k < (int)rin) ? tile_desc[(size_t)(tile - 1 - k0 - k) * 256 + threadIdx.x] : E3_LB_AGG;
#pragma unroll
			for (int k = 0; k < LB; k++) {
				while ((v[k] & E3_LB_AGG) == 0) { __nanosleep(20); v[k] = tile_desc[(size_t)(tile - 1 - k0 - k) * 256 + threadIdx.x]; }
				in_group += v[k] & E3_LB_MASK;
			}
		}
		const bool closes_group = rin == E3_SORT_GROUP - 1;
		if (closes_group && grp != 0) group_desc[(size_t)grp * 256 + threadIdx.x] = (in_group + total) | E3_LB_AGG;
		uint32_t before = 0;
		bool found = (grp == 0);
		for (int g = (int)grp - 1; g >= 0 && !found; g -= LB) {
			uint32_t v[LB];
#pragma unroll
			for (int k = 0; k < LB; k++) v[k] = (g - k >= 0) ? group_desc[(size_t)(g - k) * 256 + threadIdx.x] : E3_LB_PREFIX;
#pragma unroll
			for (int k = 0; k < LB; k++) {
				if (found) break;
				while ((v[k] & (E3_LB_AGG | E3_LB_PREFIX)) == 0) { __nanosleep(20); v[k] = group_desc[(size_t)(g - k) * 256 + threadIdx.x]; }
				before += v[k] & E3_LB_MASK;
				if (v[k] & E3_LB_PREFIX) found = true;
			}
		}
		const uint32_t excl = before + in_group;
		if (closes_group) { __threadfence(); group_desc[(size_t)grp * 256 + threadIdx.x] = (excl + total) | E3_LB_PREFIX; }
		s_gpos[threadIdx.x] = gb + excl;
		s_lstart[threadIdx.x] = ls;
	}
	__syncthreads();
	// reorder the tile in shared memory: digit runs, original order inside a run
#pragma unroll
	for (int i = 0; i < E3_SORT_ITEMS; i++) {
		if (wstart + (size_t)i * 32 + lane < n) {
			const uint32_t d = (key[i] >> shift) & mask;
			const uint32_t lp = s_lstart[d] + wcnt[w][d] + rank[i];
			s_keys[lp] = key[i];
			s_vals[lp] = val[i];
		}
	}
	__syncthreads();
	for (uint32_t i = threadIdx.x; i < tcount; i += E3_SORT_THREADS) {
		const uint32_t k = s_keys[i];
		const uint32_t d = (k >> shift) & mask;
		const uint32_t pos = s_gpos[d] + (i - s_lstart[d]);
		keys_out[pos] = k;
		vals_out[pos] = s_vals[i];
	}
}

// ---------------- instance emission (duplicateWithKeys, rasterizer_impl.cu:70-111) ----------------
// One thread per Gaussian in DEPTH order. Small footprints are written by the owning lane;
// footprints of more than 32 tiles are written cooperatively by the whole warp, so one large
// splat does not stall 31 idle lanes (the reference's serial inner loop does).
__global__ void __launch_bounds__(256) emit_instances_kernel(int P, const uint32_t* __restrict__ order,
	const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ tiles_touched,
	const uint2* __restrict__ rect, int gx, uint32_t* __restrict__ tkeys, uint32_t* __restrict__ vals, uint32_t cap)
{
	// cap: capacity of tkeys / vals. With a host-known R it equals R; on the no-sync path it is the
	// caller's estimate and instances beyond it are dropped (the host detects R > cap and redoes the step).
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
			for (uint32_t x = x0; x < x1; x++) { if (o < cap) { tkeys[o] = y * gx + x; vals[o] = idx; } o++; }
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
			if (b_off + i < cap) {
				tkeys[b_off + i] = (b_y0 + yy) * gx + (b_x0 + xx);
				vals[b_off + i] = b_idx;
			}
		}
	}
}

// identifyTileRanges, rasterizer_impl.cu:116-138, on 32-bit tile ids
__global__ void __launch_bounds__(256) tile_ranges_kernel(size_t L, const uint32_t* __restrict__ tkeys,
	uint2* __restrict__ ranges, const uint32_t* __restrict__ n_dev)
{
	if (n_dev) L = min(L, (size_t)*n_dev);
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
	uint32_t* counts, uint32_t* scan_tmp, int start_buf, int* result_buf, cudaStream_t st, const uint32_t* n_dev,
	bool all_ones_is_padding)
{
	(void)scan_tmp;
	int cur = start_buf;
	if (n > 0 && end_bit > 0) {
		RadixPlan plan;
		plan.npass = e3_radix_passes(end_bit);
		const size_t nblk = e3_sort_blocks(n);
		int lo = 0;
		for (int p = 0; p < 4; p++) { plan.shift[p] = 0; plan.bits[p] = 1; }
		for (int p = 0; p < plan.npass; p++) {
			// split the bits evenly over the passes (e.g. 12 -> 6 + 6): fewer digits per pass
			// means longer contiguous runs per tile in the scatter
			const int nb = (end_bit - lo + (plan.npass - p) - 1) / (plan.npass - p);
			plan.shift[p] = lo; plan.bits[p] = nb;
			lo += nb;
		}
		uint32_t* ghist = counts;
		uint32_t* tile_counter = counts + 4 * 256;
		uint32_t* lookback = counts + 4 * 256 + 8;
		cudaError_t e = cudaMemsetAsync(counts, 0, e3_sort_scratch_words(n) * sizeof(uint32_t), st);
		if (e != cudaSuccess) return e;
		const unsigned hgrid = (unsigned)(nblk < 1184 ? nblk : 1184);
		radix_hist_all_kernel<<<hgrid, 256, 0, st>>>(keys[cur], n, plan, ghist, n_dev);
		for (int p = 0; p < plan.npass; p++) {
			radix_onesweep_kernel<<<(unsigned)nblk, E3_SORT_THREADS, 0, st>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n,
				plan.shift[p], (1u << plan.bits[p]) - 1u, ghist + p * 256, lookback + (size_t)p * e3_sort_desc_words(n), tile_counter + p, n_dev,
				(all_ones_is_padding && p == plan.npass - 1 && plan.shift[p] + plan.bits[p] == 32) ? (int)((1u << plan.bits[p]) - 1u) : -1);
			cur ^= 1;
		}
		g_e3_launches += 1 + plan.npass;
	}
	*result_buf = cur;
	return cudaGetLastError();
}

cudaError_t e3_launch_emit_instances(int P, const uint32_t* order, const uint32_t* offsets,
	const uint32_t* tiles_touched, const uint2* rect, int gx, uint32_t* tkeys, uint32_t* vals, uint32_t cap,
	cudaStream_t st)
{
	emit_instances_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, order, offsets, tiles_touched, rect, gx, tkeys, vals, cap);
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_tile_ranges(size_t R, const uint32_t* tkeys, uint2* ranges, size_t T, cudaStream_t st, const uint32_t* n_dev)
{
	cudaError_t e = cudaMemsetAsync(ranges, 0, T * sizeof(uint2), st);
	if (e != cudaSuccess) return e;
	if (R > 0) { tile_ranges_kernel<<<(unsigned)((R + 255) / 256), 256, 0, st>>>(R, tkeys, ranges, n_dev); g_e3_launches += 1; }
	return cudaGetLastError();
}
