// Tile blending, forward and backward. sm_100a.
//
// Semantics follow the reference renderCUDA kernels (cuda_rasterizer/forward.cu:261-374,
// backward.cu:399-557) pixel for pixel: same front-to-back order, same skips (power > 0,
// alpha < 1/255), same stop (T*(1-alpha) < 1e-4), same `contributor` bookkeeping, and in the
// forward the same rounding sequence (explicit fma/mul/add in the reference's SASS contraction
// pattern), so out_color / final_T / n_contrib are bit-identical to the reference build.
//
// What differs is the machine mapping:
//   * one CTA per 16x16 tile, each warp owns an 8x4 pixel block (better spatial coherence for
//     the warp-level votes than the reference's 16x2 rows);
//   * per-instance attributes come from ONE packed 48-byte record per Gaussian, gathered with
//     three 16-byte loads into shared memory (the reference gathers 4+8+16 bytes from three
//     arrays and fetches colours from global memory per pixel);
//   * warps whose 32 pixels are all finished skip the batch; the CTA leaves when all are;
//   * backward: the gradient terms of a (pixel, Gaussian) pair leave the lane as two 16-byte vector
//     reductions (red.global.add.v4.f32) into a packed 48-byte accumulator row when at most kDirectMax
//     pixels of the warp's block take part; otherwise the warp sums them first with a reduce-scatter
//     butterfly (14 shuffles for 9 values instead of 45) and ONE red.global per value leaves the warp
//     -- versus 9 same-address scalar atomics per (pixel, Gaussian) pair in the reference (backward.cu:523-554);
//   * backward starts at the tile's last contributing instance instead of the end of the range.
#include "common.cuh"

namespace {

#ifndef E3_BLEND_BATCH_FWD
#define E3_BLEND_BATCH_FWD 512
#endif
#ifndef E3_BLEND_BATCH_BWD
#define E3_BLEND_BATCH_BWD 512
#endif
// instances staged per barrier pair (multiples of the 256 threads). Measured at c3, serial blends per iteration:
// forward 128: 0.598, 256: 0.585, 512: 0.571 ms (it re-checks for finished tiles at every batch, and 21 % of its stall
// samples sit at the two barriers of a batch); backward 256: 1.111, 512: 1.095, 1024: 1.132 ms (no early exit; the
// warps of a tile do different amounts of work per batch and wait for each other at the barriers).
constexpr int kBatchF = E3_BLEND_BATCH_FWD, kBatchB = E3_BLEND_BATCH_BWD;
constexpr float kLog2e = 1.4426950408889634f;
// backward blend: up to this many contributing pixels of a (block, instance) pair send their gradient terms to the
// accumulator row themselves; with more, the warp sums them first. Measured at c3 (serial backward blends per iteration,
// ms): 0: 1.35, 4: 1.22, 8: 1.16, 12: 1.12, 16: 1.10, 20: 1.12, 24: 1.21, 32: 1.73 (the L2's reduction rate binds).
#ifndef E3_BWD_DIRECT_MAX
#define E3_BWD_DIRECT_MAX 16
#endif
constexpr int kDirectMax = E3_BWD_DIRECT_MAX;

__device__ __forceinline__ void pixel_of_thread(int& lx, int& ly)
{
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	lx = ((w & 1) << 3) + (lane & 7);
	ly = ((w >> 1) << 2) + (lane >> 3);
}

// Which of the tile's eight 8x4 warp blocks can an instance touch at all? A pixel receives a
// contribution only if alpha = opacity * exp(power) >= 1/255, i.e. inside the ellipse
// q(d) = conic.x dx^2 + 2 conic.y dx dy + conic.z dy^2 <= 2 tau, tau = ln(255 * opacity) (+ margin), all
// precomputed per Gaussian (preprocess.cu): (hx, hy) is the half-extent of the ellipse's bounding box.
// Two steps: the bounding box against the block, then -- for the blocks that pass -- the exact minimum of q over
// the block's rectangle of pixel centres: 0 if the centre lies inside, otherwise attained on one of the (at most
// two) faces that look at the centre, where it is a clamped one-dimensional minimisation. The second step removes
// the blocks that only a corner of the bounding box reaches (about one examined (block, instance) pair in six).
// Bit w = (row r)*2 + (column c) matches pixel_of_thread(). Both steps are conservative (margins cover the fp32
// rounding of the blend's own power evaluation; anything odd -- NaN, indefinite conic -- keeps the block), so
// skipping changes no result bit: the per-pixel test would have skipped the same pairs.
__device__ __forceinline__ uint32_t subtile_mask(float x, float y, float hx, float hy, float cx, float cy, float cz, float tau)
{
	if (!(hx >= 0.f)) return 0u;
	const float xl = x - hx, xh = x + hx, yl = y - hy, yh = y + hy;
	const uint32_t cm = ((xh >= 0.f) & (xl <= 7.f)) | (((xh >= 8.f) & (xl <= 15.f)) << 1);
	uint32_t m = 0;
#pragma unroll
	for (int r = 0; r < 4; r++)
		if ((yh >= 4.f * r) & (yl <= 4.f * r + 3.f)) m |= cm << (2 * r);
	if (m == 0u) return 0u;
	const float ry = __fdividef(-cy, cz), rx = __fdividef(-cy, cx), cy2 = cy + cy;
	const float thr = 2.002f * tau + 2e-3f;
	uint32_t out = 0;
#pragma unroll
	for (int c = 0; c < 2; c++) {
		const float X0 = 8.f * c - x, X1 = X0 + 7.f;          // pixel centre minus Gaussian centre
		const float xc = fminf(fmaxf(0.f, X0), X1);
		const float xm = fmaxf(fabsf(X0), fabsf(X1));
#pragma unroll
		for (int r = 0; r < 4; r++) {
			if (!((m >> (2 * r + c)) & 1u)) continue;
			const float Y0 = 4.f * r - y, Y1 = Y0 + 3.f;
			const float yc = fminf(fmaxf(0.f, Y0), Y1);
			const float ym = fmaxf(fabsf(Y0), fabsf(Y1));
			const float yy = fminf(fmaxf(ry * xc, Y0), Y1);     // face dx = xc
			const float xx = fminf(fmaxf(rx * yc, X0), X1);     // face dy = yc
			float q1 = xc * fmaf(cx, xc, cy2 * yy) + cz * yy * yy;
			float q2 = xx * fmaf(cx, xx, cy2 * yc) + cz * yc * yc;
			q1 = xc != 0.f ? q1 : 3.0e38f;
			q2 = yc != 0.f ? q2 : 3.0e38f;
			float q = fminf(q1, q2);
			q = (xc == 0.f && yc == 0.f) ? 0.f : q;
			// the blend evaluates the power with a few ulp of its largest term
			const float slack = 4e-6f * fmaf(cx, xm * xm, cz * ym * ym);
			if (!(q > thr + slack)) out |= 1u << (2 * r + c);
		}
	}
	return out;
}

// Shared-memory reads through an explicit 32-bit shared address, computed once per thread. With
// plain array indexing the compiler re-derives the CTA's shared window base (S2UR SR_CgaCtaId +
// UMOV + ULEA) at every use inside the inner loop; these keep it in one register.
__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
	uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
	asm volatile("" : "+r"(a));     // opaque: keep it in a register instead of re-materialising it
	return a;
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr)
{
	float4 v;
	asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
	return v;
}
__device__ __forceinline__ float lds_f(uint32_t addr)
{
	float v;
	asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
	return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr)
{
	uint32_t v;
	asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
	return v;
}
__device__ __forceinline__ uint32_t lds_u8(uint32_t addr)
{
	uint32_t v;
	asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
	return v;
}

// position of the most significant set bit (FLO.U32; m != 0)
__device__ __forceinline__ int top_bit(uint32_t m)
{
	int p;
	asm("bfind.u32 %0, %1;" : "=r"(p) : "r"(m));
	return p;
}

// 1u << p without a constant register (BMSK)
__device__ __forceinline__ uint32_t bit_at(int p)
{
	uint32_t b;
	asm("bmsk.clamp.b32 %0, %1, 1;" : "=r"(b) : "r"(p));
	return b;
}

// 16-byte vector reduction (the accumulator rows are 48 bytes, 16-byte aligned)
__device__ __forceinline__ void red_add4(float* p, float x, float y, float z, float w)
{
	asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" :: "l"(p), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

// forward.cu:335 with the reference's contraction
__device__ __forceinline__ float eval_power(float dx, float dy, float cx, float cy, float cz)
{
	const float sum = __fmaf_rn(dx, __fmul_rn(cx, dx), __fmul_rn(__fmul_rn(cz, dy), dy));
	return __fmaf_rn(sum, -0.5f, -__fmul_rn(__fmul_rn(cy, dx), dy));
}

// EXACT: the accurate expf of the reference build (forward.cu:336) -- the drop-in path, whose image, final_T and
// n_contrib are bit-identical to the reference's. !EXACT (fused path, held to the 1e-5 image tolerance instead):
// ex2.approx of power * log2(e), the very expression the backward blend evaluates, so that the backward sees the
// alphas the forward blended with; 6 instructions fewer per examined pair.
template <int C, bool EXACT>
__global__ void __launch_bounds__(256) render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
	int gx, const uint32_t* __restrict__ point_list, int W, int H, const float4* __restrict__ rec,
	const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
	float* __restrict__ out_color)
{
	// {x, y, conic.x, conic.y} and {conic.z, opacity, c0, c1} of instance j at sAB[2j], sAB[2j+1]:
	// one address computation, two LDS.128 at fixed offsets
	__shared__ float4 sAB[kBatchF * 2];
	__shared__ float sC[kBatchF];    // c2
	__shared__ uint8_t sMask[kBatchF];
	int lx, ly;
	pixel_of_thread(lx, ly);
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t tile = tile_order[blockIdx.x];       // longest tiles are scheduled first
	const int tx = tile % gx, ty = tile / gx;
	const int px = tx * E3_TILE_X + lx, py = ty * E3_TILE_Y + ly;
	const bool inside = px < W && py < H;
	const uint32_t pix_id = (uint32_t)W * py + px;
	const float pixfx = (float)px, pixfy = (float)py;
	const float tile_x0 = (float)(tx * E3_TILE_X), tile_y0 = (float)(ty * E3_TILE_Y);
	const uint2 range = ranges[tile];
	const int total = (int)(range.y - range.x);
	// T < 0 marks a finished pixel (the reference's `done`): its magnitude is the transmittance it stopped with, and
	// T (1 - alpha) < 0 fails the blend test from then on without a separate flag. Pixels outside the image start finished.
	float T = inside ? 1.0f : -1.0f;
	uint32_t last_contributor = 0;
	const uint32_t aAB = smem_addr(sAB), aC = smem_addr(sC), aMask = smem_addr(sMask);
	float Cacc[C];
#pragma unroll
	for (int c = 0; c < C; c++) Cacc[c] = 0.f;

	for (int b0 = 0; b0 < total; b0 += kBatchF) {
		if (__syncthreads_count(T < 0.f) == E3_TILE_PIX) break;
#pragma unroll
		for (int l = threadIdx.x; l < kBatchF; l += 256) {
			const int k = b0 + l;
			if (k < total) {
				const uint32_t id = point_list[range.x + k];
				const float4* r = rec + 3 * (size_t)id;
				const float4 a = __ldg(r);
				const float4 e = __ldg(r + 2);     // c2, hx, hy, -
				sAB[2 * l] = a;
				if (C == 3) sC[l] = e.x;
				const float4 b = __ldg(r + 1);
				sAB[2 * l + 1] = b;
				sMask[l] = (uint8_t)subtile_mask(a.x - tile_x0, a.y - tile_y0, e.y, e.z, a.z, a.w, b.x, e.w);
			}
		}
		__syncthreads();
		if (__all_sync(E3_FULL, T < 0.f)) continue;    // this warp's 32 pixels are finished
		const int nb = min(kBatchF, total - b0);
		for (int c0 = 0; c0 < nb; c0 += 32) {
			// lane l looks at instance c0 + 31 - l: the instances of the chunk that reach this warp's block, first
			// instance in the top bit, so that one count-leading-zeros yields the next one
			const int jj = c0 + 31 - lane;
			uint32_t m = __ballot_sync(E3_FULL, jj < nb && ((lds_u8(aMask + jj) >> w) & 1u));
			const uint32_t cAB = aAB + 32 * (c0 + 31), cC = aC + 4 * (c0 + 31);
			const uint32_t contrib_no = (uint32_t)(b0 + c0 + 32);
			while (m) {
				const int p = top_bit(m);                // instance c0 + 31 - p
				m ^= bit_at(p);
				const float4 a = lds_f4(cAB - 32 * p);
				const float4 b = lds_f4(cAB - 32 * p + 16);
				const float dx = __fadd_rn(a.x, -pixfx), dy = __fadd_rn(a.y, -pixfy);
				const float power = eval_power(dx, dy, a.z, a.w, b.x);
				// branch-free: every lane evaluates, selects keep the state of lanes that skip
				float G;
				if (EXACT) G = expf(power);
				else asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(power * kLog2e));
				const float alpha = fminf(0.99f, __fmul_rn(b.y, G));
				const bool contrib = power <= 0.0f && alpha >= 1.0f / 255.0f;
				const float test_T = __fmul_rn(T, __fadd_rn(1.0f, -alpha));
				const bool blend = contrib && test_T >= 0.0001f;
				const float wgt = __fmul_rn(alpha, b.z);
				Cacc[0] = blend ? __fmaf_rn(T, wgt, Cacc[0]) : Cacc[0];
				if (C == 3) {
					Cacc[1] = blend ? __fmaf_rn(T, __fmul_rn(alpha, b.w), Cacc[1]) : Cacc[1];
					Cacc[2] = blend ? __fmaf_rn(T, __fmul_rn(alpha, lds_f(cC - 4 * p)), Cacc[2]) : Cacc[2];
				}
				last_contributor = blend ? contrib_no - (uint32_t)p : last_contributor;
				T = blend ? test_T : T;
				T = (contrib && !blend) ? -fabsf(T) : T;     // finished (forward.cu:348-352)
			}
			if (__all_sync(E3_FULL, T < 0.f)) break;
		}
	}
	T = fabsf(T);
	if (inside) {
		final_T[pix_id] = T;
		n_contrib[pix_id] = last_contributor;
		const size_t HW = (size_t)H * W;
#pragma unroll
		for (int c = 0; c < C; c++) out_color[c * HW + pix_id] = __fmaf_rn(__ldg(bg + c), T, Cacc[c]);
	}
}

// Sum v[0..7] over the 32 lanes with a reduce-scatter butterfly. On return lane L with
// (L & 3) == 0 holds the full sum of slot ((L>>4)&1)*4 + ((L>>3)&1)*2 + ((L>>2)&1) in v[0].
__device__ __forceinline__ void warp_reduce8(float (&v)[8], int lane)
{
	{
		const bool hi = lane & 16;
#pragma unroll
		for (int k = 0; k < 4; k++) {
			const float send = hi ? v[k] : v[k + 4];
			const float keep = hi ? v[k + 4] : v[k];
			v[k] = keep + __shfl_xor_sync(E3_FULL, send, 16);
		}
	}
	{
		const bool hi = lane & 8;
#pragma unroll
		for (int k = 0; k < 2; k++) {
			const float send = hi ? v[k] : v[k + 2];
			const float keep = hi ? v[k + 2] : v[k];
			v[k] = keep + __shfl_xor_sync(E3_FULL, send, 8);
		}
	}
	{
		const bool hi = lane & 4;
		const float send = hi ? v[0] : v[1];
		const float keep = hi ? v[1] : v[0];
		v[0] = keep + __shfl_xor_sync(E3_FULL, send, 4);
	}
	v[0] += __shfl_xor_sync(E3_FULL, v[0], 2);
	v[0] += __shfl_xor_sync(E3_FULL, v[0], 1);
}

// Same for 16 slots: on return lane L with (L & 1) == 0 holds the full sum of slot
// ((L>>4)&1)*8 + ((L>>3)&1)*4 + ((L>>2)&1)*2 + ((L>>1)&1) in v[0]. 16 shuffles.
__device__ __forceinline__ void warp_reduce16(float (&v)[16], int lane)
{
#pragma unroll
	for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
		const bool hi = lane & bit;
#pragma unroll
		for (int k = 0; k < 8; k++) {
			if (k < half) {
				const float send = hi ? v[k] : v[k + half];
				const float keep = hi ? v[k + half] : v[k];
				v[k] = keep + __shfl_xor_sync(E3_FULL, send, bit);
			}
		}
	}
	v[0] += __shfl_xor_sync(E3_FULL, v[0], 1);
}

// DUAL (3 channels only): a second seed dL_dpixels2 rides along, of which only the screen-space
// gradient is wanted (accumulator slots 9, 10). Used when the intensity view and the first event view of
// an iteration are the same camera and are rendered once: dL_dpixels is then the SUM of the two loss
// terms' seeds (all parameter gradients are linear in it), dL_dpixels2 the intensity term alone, whose
// screen-space gradient is what the reference's densification statistics accumulate (train.py:145,319-320).
// EXACT (drop-in path): G and alpha by the forward's own instruction sequence (accurate expf, the same rounded
// product), so the backward takes exactly the forward's -- and the reference's -- alpha >= 1/255 decisions. A loss
// like the event loss (log of the rendered luma) puts seeds of 1e6 times the normal size on nearly black pixels;
// there a contribution that one pass blends and the other skips on the last bit of alpha decides whole gradient rows.
// !EXACT (fused path): ex2.approx in both passes, equally consistent with each other.
template <int C, bool DUAL, bool EXACT>
__global__ void __launch_bounds__(256) render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
	int gx, const uint32_t* __restrict__ point_list, int W, int H, const float* __restrict__ bg,
	const float4* __restrict__ rec, const float* __restrict__ final_Ts,
	const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels, const float* __restrict__ dL_dpixels2,
	float* __restrict__ acc)
{
	__shared__ float4 sAB[kBatchB * 2];
	__shared__ float sC[kBatchB];
	__shared__ uint32_t sId[kBatchB];
	__shared__ uint8_t sMask[kBatchB];
	__shared__ int s_max[8];
	int lx, ly;
	pixel_of_thread(lx, ly);
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const uint32_t tile = tile_order[blockIdx.x];
	const int tx = tile % gx, ty = tile / gx;
	const int px = tx * E3_TILE_X + lx, py = ty * E3_TILE_Y + ly;
	const bool inside = px < W && py < H;
	const uint32_t pix_id = (uint32_t)W * py + px;
	const float pixfx = (float)px, pixfy = (float)py;
	const float tile_x0 = (float)(tx * E3_TILE_X), tile_y0 = (float)(ty * E3_TILE_Y);
	const uint2 range = ranges[tile];
	const float T_final = inside ? final_Ts[pix_id] : 0.f;
	float T = T_final;
	const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;

	// nothing behind the tile's deepest contributor can receive a gradient: start there
	int wmax = last_contributor;
#pragma unroll
	for (int o = 16; o; o >>= 1) wmax = max(wmax, __shfl_xor_sync(E3_FULL, wmax, o));
	if (lane == 0) s_max[w] = wmax;
	__syncthreads();
	int total = 0;
#pragma unroll
	for (int i = 0; i < 8; i++) total = max(total, s_max[i]);

	// Bp = sum_c dL/dpixel_c * (colour seen behind the current instance, normalised by the transmittance in
	// front of what lies behind): the background to start with. With pc = sum_c dL/dpixel_c colour_c this is
	// backward.cu:505-534 (accum_rec recurrence + background term) pre-multiplied with dL/dpixel:
	//   dL/dalpha_i = T_i (pc_i - Bp);   Bp <- Bp + alpha_i (pc_i - Bp)
	float dL_dpixel[C], dL_dpixel2[C];
	const size_t HW = (size_t)H * W;
	float Bp = 0.f, Bp2 = 0.f;
#pragma unroll
	for (int c = 0; c < C; c++) {
		dL_dpixel[c] = inside ? dL_dpixels[c * HW + pix_id] : 0.f;
		dL_dpixel2[c] = (DUAL && inside) ? dL_dpixels2[c * HW + pix_id] : 0.f;
		const float b = __ldg(bg + c);
		Bp += b * dL_dpixel[c];
		Bp2 += b * dL_dpixel2[c];
	}
	const uint32_t aAB = smem_addr(sAB), aC = smem_addr(sC), aMask = smem_addr(sMask), aId = smem_addr(sId);
	// after warp_reduce8 lane L with (L & 3) == 0 holds slot my_slot of the accumulator row
	const int my_slot = DUAL ? ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1)
		: ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);

	// instance index within the tile's list, walked from total-1 down to 0; an instance with
	// 0-based index i has contributor number i (reference: `contributor--` before the test,
	// valid iff contributor < last_contributor).
	for (int b0 = 0; b0 < total; b0 += kBatchB) {
		__syncthreads();
#pragma unroll
		for (int l = threadIdx.x; l < kBatchB; l += 256) {
			const int k = b0 + l;
			if (k < total) {
				const uint32_t id = point_list[range.x + (total - 1 - k)];
				const float4* r = rec + 3 * (size_t)id;
				const float4 a = __ldg(r);
				const float4 e = __ldg(r + 2);
				sId[l] = id;
				const float4 b = __ldg(r + 1);
				sAB[2 * l] = a;
				sAB[2 * l + 1] = b;
				if (C == 3) sC[l] = e.x;
				sMask[l] = (uint8_t)subtile_mask(a.x - tile_x0, a.y - tile_y0, e.y, e.z, a.z, a.w, b.x, e.w);
			}
		}
		__syncthreads();
		const int nb = min(kBatchB, total - b0);
		// skip the part of the batch that lies behind this warp's deepest contributor
		const int j0 = max(0, (total - b0) - wmax);
		const int jlim = total - 1 - b0 - last_contributor;
		for (int c0 = (j0 & ~31); c0 < nb; c0 += 32) {
		// lane l looks at instance c0 + 31 - l: first instance of the chunk in the top bit, so that one
		// count-leading-zeros yields the next one
		const int jj = c0 + 31 - lane;
		uint32_t m = __ballot_sync(E3_FULL, jj < nb && jj >= j0 && ((lds_u8(aMask + jj) >> w) & 1u));
		const uint32_t cAB = aAB + 32 * (c0 + 31), cC = aC + 4 * (c0 + 31), cId = aId + 4 * (c0 + 31);
		const int plim = c0 + 31 - jlim;
		while (m) {
			const int p = top_bit(m);                // instance j = c0 + 31 - p
			m ^= bit_at(p);
			const float4 a = lds_f4(cAB - 32 * p);
			const float4 b = lds_f4(cAB - 32 * p + 16);
			const float dx = a.x - pixfx, dy = a.y - pixfy;
			// the forward's own rounding of the power: for a thin, oblique ellipse its terms cancel to a small
			// difference, and another evaluation order would give another G than the one the forward blended with
			const float power = eval_power(dx, dy, a.z, a.w, b.x);
			// ex2.approx of power * log2(e): within 5e-7 relative of the forward's accurate expf, far inside
			// the gradient tolerance (1e-4), 6 instructions shorter per examined pair
			float G;
			if (EXACT) G = expf(power);
			else asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(G) : "f"(power * kLog2e));
			float alpha = fminf(0.99f, __fmul_rn(b.y, G));
			// contributor = total - 1 - (b0 + j) < last_contributor  <=>  j > jlim
			const bool valid = (p < plim) && (power <= 0.0f) && (alpha >= 1.0f / 255.0f);
			const uint32_t bal = __ballot_sync(E3_FULL, valid);
			if (bal == 0u) continue;

			// Branch-free from here: a lane without a contribution runs the same arithmetic with
			// alpha = G = 0, which leaves its state untouched (T * 1, selects) and makes all of its
			// addends exact zeros.
			alpha = valid ? alpha : 0.f;
			G = valid ? G : 0.f;
			float v[8];
			float v8 = 0.f;
			// one reciprocal serves the division of backward.cu:503 (tolerance-checked path; rcp.approx is
			// within 1 ulp and 1 - alpha lies in [0.01, 1], no special cases)
			float ra;
			asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(1.f - alpha));
			T = T * ra;
			const float aT = alpha * T;
			float pc = dL_dpixel[0] * b.z, pc2 = dL_dpixel2[0] * b.z, c2v = 0.f;
			if (C == 3) {
				c2v = lds_f(cC - 4 * p);
				pc = fmaf(dL_dpixel[2], c2v, fmaf(dL_dpixel[1], b.w, pc));
				if (DUAL) pc2 = fmaf(dL_dpixel2[2], c2v, fmaf(dL_dpixel2[1], b.w, pc2));
			}
			const float dB = pc - Bp;
			const float g = G * (dB * T);                      // G dL/dalpha; dL/dG = opacity * g
			Bp = fmaf(alpha, dB, Bp);                          // alpha = 0 (no contribution) leaves it alone
			// What leaves the warp are sums that still lack their per-Gaussian factors (d = centre - pixel):
			//   sum g, -sum g dpower/ddx, -sum g dpower/ddy, sum g dx^2, sum g dx dy, sum g dy^2, sum alpha T dL/dpixel_c.
			// The backward preprocess, which has the opacity at hand, turns them into dL/dmean2D, dL/dconic and
			// dL/dopacity (backward.cu:538-554). The derivative of the power is formed per pixel, not from first
			// moments after the sum: along the major axis of a thin ellipse conic.x dx + conic.y dy cancels.
			const float gxv = g * dx, gyv = g * dy;
			v[0] = g;
			v[1] = fmaf(gxv, a.z, gyv * a.w);                  // g (conic.x dx + conic.y dy) = -g dpower/ddx
			v[2] = fmaf(gyv, b.x, gxv * a.w);                  // g (conic.z dy + conic.y dx) = -g dpower/ddy
			v[3] = gxv * dx;
			v[4] = gxv * dy;
			v[5] = gyv * dy;
			v[6] = aT * dL_dpixel[0];
			v[7] = 0.f;
			if (C == 3) { v[7] = aT * dL_dpixel[1]; v8 = aT * dL_dpixel[2]; }
			float u9 = 0.f, u10 = 0.f;
			if (DUAL) {
				const float dB2 = pc2 - Bp2;
				const float g2 = G * (dB2 * T);
				Bp2 = fmaf(alpha, dB2, Bp2);
				const float g2x = g2 * dx, g2y = g2 * dy;
				u9 = fmaf(g2x, a.z, g2y * a.w);
				u10 = fmaf(g2y, b.x, g2x * a.w);
			}
			float* dst = acc + (size_t)lds_u32(cId - 4 * p) * E3_ACC_STRIDE;
			// A pair that at most kDirectMax pixels of the block take part in: they add their terms themselves, two (three)
			// 16-byte vector reductions per pixel. Above that the L2's reduction rate would bind and the warp sums first.
			if (__popc(bal) <= kDirectMax) {
				if (valid) {
					red_add4(dst, v[0], v[1], v[2], v[3]);
					red_add4(dst + 4, v[4], v[5], v[6], v[7]);
					if (DUAL) red_add4(dst + 8, v8, u9, u10, 0.f);
					else if (C == 3) atomicAdd(dst + 8, v8);
				}
			} else if (DUAL) {
				float u[16];
#pragma unroll
				for (int i = 0; i < 8; i++) u[i] = v[i];
				u[8] = v8; u[9] = u9; u[10] = u10;
				u[11] = u[12] = u[13] = u[14] = u[15] = 0.f;
				warp_reduce16(u, lane);
				if ((lane & 1) == 0 && my_slot < 11) atomicAdd(dst + my_slot, u[0]);
			} else {
				warp_reduce8(v, lane);
				if (C == 3) {
#pragma unroll
					for (int o = 16; o; o >>= 1) v8 += __shfl_xor_sync(E3_FULL, v8, o);
				}
				if ((lane & 3) == 0 && (C == 3 || my_slot < 7)) atomicAdd(dst + my_slot, v[0]);
				if (C == 3 && lane == 1) atomicAdd(dst + 8, v8);
			}
		}
		}
	}
}

// Approximate longest-processing-time-first schedule: tiles are bucketed by list length (64
// classes relative to the longest list) and emitted longest class first, so that the last wave of
// CTAs of the blend kernels is made of short tiles instead of whatever the raster order leaves.
__global__ void __launch_bounds__(1024) tile_order_kernel(const uint2* __restrict__ ranges, int T, uint32_t* __restrict__ order)
{
	__shared__ uint32_t s_max, cnt[64], start[64];
	if (threadIdx.x == 0) s_max = 0;
	if (threadIdx.x < 64) cnt[threadIdx.x] = 0;
	__syncthreads();
	uint32_t m = 0;
	for (int t = threadIdx.x; t < T; t += blockDim.x) { const uint2 r = ranges[t]; m = max(m, r.y - r.x); }
	for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(E3_FULL, m, o));
	if ((threadIdx.x & 31) == 0) atomicMax(&s_max, m);
	__syncthreads();
	const uint32_t mx = s_max + 1;
	for (int t = threadIdx.x; t < T; t += blockDim.x) {
		const uint2 r = ranges[t];
		const uint32_t b = 63u - min(63u, (uint32_t)(((uint64_t)(r.y - r.x) * 64u) / mx));
		atomicAdd(&cnt[b], 1u);
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		uint32_t run = 0;
		for (int b = 0; b < 64; b++) { start[b] = run; run += cnt[b]; cnt[b] = 0; }
	}
	__syncthreads();
	for (int t = threadIdx.x; t < T; t += blockDim.x) {
		const uint2 r = ranges[t];
		const uint32_t b = 63u - min(63u, (uint32_t)(((uint64_t)(r.y - r.x) * 64u) / mx));
		order[start[b] + atomicAdd(&cnt[b], 1u)] = (uint32_t)t;
	}
}

}  // namespace

cudaError_t e3_launch_tile_order(const uint2* ranges, int T, uint32_t* order, cudaStream_t st)
{
	tile_order_kernel<<<1, 1024, 0, st>>>(ranges, T, order);
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_render_forward(const E3Frame& f, const uint2* ranges, const uint32_t* tile_order, const uint32_t* point_list,
	const float4* rec, const float* bg, float* final_T, uint32_t* n_contrib, float* out_color, cudaStream_t st, bool exact)
{
	const unsigned grid = (unsigned)(f.gx * f.gy);
#define FWD(CH, EX) render_fwd_kernel<CH, EX><<<grid, 256, 0, st>>>(ranges, tile_order, f.gx, point_list, f.W, f.H, rec, bg, final_T, n_contrib, out_color)
	if (f.C == 1) { if (exact) FWD(1, true); else FWD(1, false); }
	else { if (exact) FWD(3, true); else FWD(3, false); }
#undef FWD
	g_e3_launches += 1;
	return cudaGetLastError();
}

cudaError_t e3_launch_render_backward(const E3Frame& f, const uint2* ranges, const uint32_t* tile_order, const uint32_t* point_list,
	const float4* rec, const float* bg, const float* final_T, const uint32_t* n_contrib, const float* dL_dpix,
	float* acc, cudaStream_t st, const float* dL_dpix2, bool exact)
{
	const unsigned grid = (unsigned)(f.gx * f.gy);
	if (dL_dpix2 && f.C != 3) return cudaErrorInvalidValue;
#define BWD(CH, DU, EX) render_bwd_kernel<CH, DU, EX><<<grid, 256, 0, st>>>(ranges, tile_order, f.gx, point_list, f.W, f.H, bg, rec, final_T, n_contrib, dL_dpix, dL_dpix2, acc)
	if (dL_dpix2) { if (exact) return cudaErrorInvalidValue; BWD(3, true, false); }
	else if (f.C == 1) { if (exact) BWD(1, false, true); else BWD(1, false, false); }
	else { if (exact) BWD(3, false, true); else BWD(3, false, false); }
#undef BWD
	g_e3_launches += 1;
	return cudaGetLastError();
}
