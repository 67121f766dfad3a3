// Page kernels: dense map expansion (A2) fused with the page remap (A3/A4).
//
// Reference: Grid.points("full") origami/core/dewarp.py:514-526 and Dewarper.dewarped /
// dewarp_image origami/core/dewarp.py:608-627.  HBM-bound: 9 B written + ~1 B read per
// dewarped pixel; no tensor cores.
//
// Layout: one CTA = 64 x 4 threads covers a 256 x 32 pixel tile of the dewarped page.  A
// warp covers 128 consecutive pixels of a row; a lane owns two pixel pairs placed so that
// every warp-wide float4 map store writes 512 contiguous bytes (full 32-byte sectors) and
// every 16-bit image store 64 contiguous bytes.  A thread walks 8 consecutive rows, so the
// x-axis interpolation state and the horizontally interpolated sparse rows stay in
// registers; they are refreshed only when the row crosses into the next sparse cell, which
// is uniform across the CTA.  The y-axis state of the tile's 32 rows is computed once per
// CTA into shared memory.  Source taps are gathered through the read-only path with 32-bit
// offsets; the smooth warp makes neighbouring threads hit the same 128-byte lines.
#include <mutex>
#include <stdlib.h>

#include "og_common.cuh"
#include "og_internal.cuh"
#include <type_traits>

#ifndef OG_PAGE_MINB
#define OG_PAGE_MINB 3
#endif
#ifndef OG_PAGE_STAGE
#define OG_PAGE_STAGE 0
#endif
#ifndef OG_PAGE_PIPELINED
#define OG_PAGE_PIPELINED 1
#endif

namespace {

constexpr int TPX = 4;    // pixels per thread along x
constexpr int TBX = 64;   // threads along x
#ifndef OG_TBY
#define OG_TBY 4
#endif
#ifndef OG_TROWS
#define OG_TROWS 8
#endif
constexpr int TBY = OG_TBY;      // thread rows
constexpr int TROWS = OG_TROWS;  // rows walked per thread
constexpr int TILE_W = TPX * TBX;
constexpr int TILE_H = TBY * TROWS;

struct PageParams {
	const uint8_t *src;
	int H, W;
	int64_t src_pitch;
	const float2 *grid;
	int gh, gw, cell;
	uint8_t *dst;
	int64_t dst_pitch;
	float *map_out;
	float scale_x, scale_y;
	cudaTextureObject_t tex;  // OG_PAGE_TEX experiment: the source as a point-sampled, zero-bordered 2-D texture (0: none)
};

// image bytes of a thread's two pixel pairs (16-bit stores when aligned and complete)
__device__ __forceinline__ void store_quad(uint8_t *orow, const int (&v)[TPX], int na, int nb, bool img_al) {
	if (img_al && na == 2) __stcs(reinterpret_cast<unsigned short *>(orow), (unsigned short)(v[0] | (v[1] << 8)));
	else {
		if (na > 0) orow[0] = (uint8_t)v[0];
		if (na > 1) orow[1] = (uint8_t)v[1];
	}
	if (img_al && nb == 2) __stcs(reinterpret_cast<unsigned short *>(orow + 64), (unsigned short)(v[2] | (v[3] << 8)));
	else {
		if (nb > 0) orow[64] = (uint8_t)v[2];
		if (nb > 1) orow[65] = (uint8_t)v[3];
	}
}

struct RowAxis {
	int r0, r1;
	float b0, b1;
};

template <bool LINEAR, bool SCALED, bool WRITE_MAP, bool WRITE_IMG>
__global__ void __launch_bounds__(TBX *TBY, OG_PAGE_MINB) og_page_kernel(const __grid_constant__ PageParams p) {
	__shared__ RowAxis s_rows[TILE_H];
	const int Wd = p.gw * p.cell, Hd = p.gh * p.cell;
	const int ytile = blockIdx.y * TILE_H;
	const double inv_cell = 1.0 / (double)p.cell;
	const int tid = threadIdx.y * TBX + threadIdx.x;
	if (tid < TILE_H) {
		// y axis of cv2.resize for the 32 rows of this tile, computed once per CTA
		og_axis ay = og_resize_axis(min(ytile + tid, Hd - 1), inv_cell);
		RowAxis ra;
		ra.r0 = min(max(ay.i, 0), p.gh - 1);
		ra.r1 = min(max(ay.i + 1, 0), p.gh - 1);
		ra.b1 = ay.f;
		ra.b0 = __fsub_rn(1.f, ay.f);
		s_rows[tid] = ra;
	}
	__syncthreads();
	// A warp covers 128 consecutive pixels of a row.  Lane l owns the pixel pairs
	// {2l, 2l+1} and {64+2l, 64+2l+1} of that segment, so that each of its two float4 map
	// stores lands in a fully written, warp-contiguous 512-byte run (a thread writing its own
	// 32 bytes as two 16-byte halves costs 45 % of the write bandwidth, measured).
	const int lane = threadIdx.x & 31;
	const int xw = (blockIdx.x * TBX + (threadIdx.x & ~31)) * TPX;
	const int xa = xw + 2 * lane, xb = xa + 64;
	const int ybeg = ytile + threadIdx.y * TROWS;
	if (xa >= Wd || ybeg >= Hd) return;
	const int gw = p.gw;
	const float2 *__restrict__ grid = p.grid;

	int sx[TPX], sx1[TPX];
	float a0[TPX], a1[TPX];
#pragma unroll
	for (int k = 0; k < TPX; ++k) {
		const int x = ((k & 2) ? xb : xa) + (k & 1);
		og_axis ax = og_resize_axis_x(min(x, Wd - 1), inv_cell, gw);
		sx[k] = ax.i;
		sx1[k] = min(ax.i + 1, gw - 1);
		a1[k] = ax.f;
		a0[k] = __fsub_rn(1.f, ax.f);
	}
	// pixels of each pair that exist (Wd need not be a multiple of the tile)
	const int na = min(2, Wd - xa), nb = max(0, min(2, Wd - xb));
	const bool map_al = WRITE_MAP && (Wd & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.map_out) & 15) == 0);
	const bool img_al = WRITE_IMG && (p.dst_pitch & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.dst) & 1) == 0);
	const uint8_t *__restrict__ src = p.src;
	const int H = p.H, W = p.W, spitch = (int)p.src_pitch;
	const float scx = p.scale_x, scy = p.scale_y;

	int cur_r0 = -1, cur_r1 = -1;
	float2 h0[TPX], h1[TPX];
	float *mrow = WRITE_MAP ? p.map_out + ((size_t)ybeg * Wd + xa) * 2 : nullptr;
	uint8_t *orow = WRITE_IMG ? p.dst + (int64_t)ybeg * p.dst_pitch + xa : nullptr;

	const int nrows = min(TROWS, Hd - ybeg);
	for (int yy = 0; yy < nrows; ++yy) {
		const RowAxis ra = s_rows[threadIdx.y * TROWS + yy];
		if (ra.r0 != cur_r0 || ra.r1 != cur_r1) {
			cur_r0 = ra.r0;
			cur_r1 = ra.r1;
			const float2 *g0 = grid + ra.r0 * gw;
			const float2 *g1 = grid + ra.r1 * gw;
#pragma unroll
			for (int k = 0; k < TPX; ++k) {
				float2 s00 = __ldg(g0 + sx[k]), s01 = __ldg(g0 + sx1[k]);
				float2 s10 = __ldg(g1 + sx[k]), s11 = __ldg(g1 + sx1[k]);
				h0[k].x = og_lerp_cv(s00.x, s01.x, a0[k], a1[k]);
				h0[k].y = og_lerp_cv(s00.y, s01.y, a0[k], a1[k]);
				h1[k].x = og_lerp_cv(s10.x, s11.x, a0[k], a1[k]);
				h1[k].y = og_lerp_cv(s10.y, s11.y, a0[k], a1[k]);
			}
		}
		float2 d[TPX];
#pragma unroll
		for (int k = 0; k < TPX; ++k) {
			d[k].x = og_lerp_cv(h0[k].x, h1[k].x, ra.b0, ra.b1);
			d[k].y = og_lerp_cv(h0[k].y, h1[k].y, ra.b0, ra.b1);
		}
		if (WRITE_MAP) {
			if (map_al && na == 2) __stcs(reinterpret_cast<float4 *>(mrow), make_float4(d[0].x, d[0].y, d[1].x, d[1].y));
			else {
				if (na > 0) __stcs(reinterpret_cast<float2 *>(mrow), d[0]);
				if (na > 1) __stcs(reinterpret_cast<float2 *>(mrow) + 1, d[1]);
			}
			if (map_al && nb == 2) __stcs(reinterpret_cast<float4 *>(mrow + 128), make_float4(d[2].x, d[2].y, d[3].x, d[3].y));
			else {
				if (nb > 0) __stcs(reinterpret_cast<float2 *>(mrow + 128), d[2]);
				if (nb > 1) __stcs(reinterpret_cast<float2 *>(mrow + 128) + 1, d[3]);
			}
			mrow += (size_t)Wd * 2;
		}
		if (WRITE_IMG) {
			float mx[TPX], my[TPX];
#pragma unroll
			for (int k = 0; k < TPX; ++k) {
				mx[k] = SCALED ? __fmul_rn(d[k].x, scx) : d[k].x;
				my[k] = SCALED ? __fmul_rn(d[k].y, scy) : d[k].y;
			}
			{
				int v[TPX];
				if (LINEAR) {
					og_sample_linear_batch<TPX>(src, H, W, spitch, mx, my, v);
				} else {
#pragma unroll
					for (int k = 0; k < TPX; ++k) v[k] = og_sample_nearest_fast(src, H, W, spitch, mx[k], my[k]);
				}
				store_quad(orow, v, na, nb, img_al);
				orow += p.dst_pitch;
			}
		}
	}
}

// ---------------------------------------------------------------------------------------
// Staged variant (LINEAR image output): the source window a tile samples is copied to shared
// memory with 16-byte cp.async transfers (zero-filled outside the image, which is exactly
// cv2.remap's BORDER_CONSTANT 0 per tap), so the four taps of a pixel are shared-memory loads
// instead of a dependent global gather.
//
// Pass A walks the thread's rows once: it evaluates the map (and streams it out when the map is
// wanted) and tracks the bounding box of the sampling coordinates; a block reduction gives the
// window.  After the copy, pass B evaluates the map again (6 flops per pixel, cheaper than
// keeping 64 coordinates in registers), quantises and blends from shared memory.  A pixel
// outside the window (cannot happen: the box is exact) or a tile whose window exceeds the
// shared-memory budget takes the global-memory path.
// ---------------------------------------------------------------------------------------
constexpr int WIN_BYTES = 20 * 1024;

__device__ __forceinline__ void og_cp_async16(void *smem_dst, const void *gmem_src, int src_bytes) {
	const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "r"(src_bytes) : "memory");
}

template <bool SCALED, bool WRITE_MAP>
__global__ void __launch_bounds__(TBX *TBY, OG_PAGE_MINB) og_page_staged_kernel(const __grid_constant__ PageParams p) {
	__shared__ RowAxis s_rows[TILE_H];
	__shared__ int s_box[4];
	__shared__ __align__(16) uint8_t s_win[WIN_BYTES];
	const int Wd = p.gw * p.cell, Hd = p.gh * p.cell;
	const int ytile = blockIdx.y * TILE_H;
	const double inv_cell = 1.0 / (double)p.cell;
	const int tid = threadIdx.y * TBX + threadIdx.x;
	if (tid < TILE_H) {
		og_axis ay = og_resize_axis(min(ytile + tid, Hd - 1), inv_cell);
		RowAxis ra;
		ra.r0 = min(max(ay.i, 0), p.gh - 1);
		ra.r1 = min(max(ay.i + 1, 0), p.gh - 1);
		ra.b1 = ay.f;
		ra.b0 = __fsub_rn(1.f, ay.f);
		s_rows[tid] = ra;
	}
	if (tid < 4) s_box[tid] = (tid < 2) ? INT_MAX : INT_MIN;
	__syncthreads();

	const int lane = threadIdx.x & 31;
	const int xw = (blockIdx.x * TBX + (threadIdx.x & ~31)) * TPX;
	const int xa = xw + 2 * lane, xb = xa + 64;
	const int ybeg = ytile + threadIdx.y * TROWS;
	const bool active = xa < Wd && ybeg < Hd;
	const int gw = p.gw;
	const float2 *__restrict__ grid = p.grid;

	int sx[TPX], sx1[TPX];
	float a0[TPX], a1[TPX];
#pragma unroll
	for (int k = 0; k < TPX; ++k) {
		const int x = ((k & 2) ? xb : xa) + (k & 1);
		og_axis ax = og_resize_axis_x(min(x, Wd - 1), inv_cell, gw);
		sx[k] = ax.i;
		sx1[k] = min(ax.i + 1, gw - 1);
		a1[k] = ax.f;
		a0[k] = __fsub_rn(1.f, ax.f);
	}
	const int na = active ? min(2, Wd - xa) : 0, nb = active ? max(0, min(2, Wd - xb)) : 0;
	const bool map_al = WRITE_MAP && (Wd & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.map_out) & 15) == 0);
	const bool img_al = (p.dst_pitch & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.dst) & 1) == 0);
	const uint8_t *__restrict__ src = p.src;
	const int H = p.H, W = p.W, spitch = (int)p.src_pitch;
	const float scx = p.scale_x, scy = p.scale_y;
	const int nrows = active ? min(TROWS, Hd - ybeg) : 0;

	int cur_r0 = -1, cur_r1 = -1;
	float2 h0[TPX], h1[TPX];
	auto eval_row = [&](int yy, float2 (&d)[TPX]) {
		const RowAxis ra = s_rows[threadIdx.y * TROWS + yy];
		if (ra.r0 != cur_r0 || ra.r1 != cur_r1) {
			cur_r0 = ra.r0;
			cur_r1 = ra.r1;
			const float2 *g0 = grid + ra.r0 * gw;
			const float2 *g1 = grid + ra.r1 * gw;
#pragma unroll
			for (int k = 0; k < TPX; ++k) {
				float2 s00 = __ldg(g0 + sx[k]), s01 = __ldg(g0 + sx1[k]);
				float2 s10 = __ldg(g1 + sx[k]), s11 = __ldg(g1 + sx1[k]);
				h0[k].x = og_lerp_cv(s00.x, s01.x, a0[k], a1[k]);
				h0[k].y = og_lerp_cv(s00.y, s01.y, a0[k], a1[k]);
				h1[k].x = og_lerp_cv(s10.x, s11.x, a0[k], a1[k]);
				h1[k].y = og_lerp_cv(s10.y, s11.y, a0[k], a1[k]);
			}
		}
#pragma unroll
		for (int k = 0; k < TPX; ++k) {
			d[k].x = og_lerp_cv(h0[k].x, h1[k].x, ra.b0, ra.b1);
			d[k].y = og_lerp_cv(h0[k].y, h1[k].y, ra.b0, ra.b1);
		}
	};

	// ---- pass A: the map itself + bounding box of the sampling coordinates
	float bx0 = 3.0e38f, by0 = 3.0e38f, bx1 = -3.0e38f, by1 = -3.0e38f;
	{
		float *mrow = WRITE_MAP ? p.map_out + ((size_t)ybeg * Wd + xa) * 2 : nullptr;
		for (int yy = 0; yy < nrows; ++yy) {
			float2 d[TPX];
			eval_row(yy, d);
			if (WRITE_MAP) {
				if (map_al && na == 2) __stcs(reinterpret_cast<float4 *>(mrow), make_float4(d[0].x, d[0].y, d[1].x, d[1].y));
				else {
					if (na > 0) __stcs(reinterpret_cast<float2 *>(mrow), d[0]);
					if (na > 1) __stcs(reinterpret_cast<float2 *>(mrow) + 1, d[1]);
				}
				if (map_al && nb == 2) __stcs(reinterpret_cast<float4 *>(mrow + 128), make_float4(d[2].x, d[2].y, d[3].x, d[3].y));
				else {
					if (nb > 0) __stcs(reinterpret_cast<float2 *>(mrow + 128), d[2]);
					if (nb > 1) __stcs(reinterpret_cast<float2 *>(mrow + 128) + 1, d[3]);
				}
				mrow += (size_t)Wd * 2;
			}
#pragma unroll
			for (int k = 0; k < TPX; ++k) {  // columns beyond Wd repeat column Wd-1: harmless
				const float mx = SCALED ? __fmul_rn(d[k].x, scx) : d[k].x;
				const float my = SCALED ? __fmul_rn(d[k].y, scy) : d[k].y;
				bx0 = fminf(bx0, mx); bx1 = fmaxf(bx1, mx);
				by0 = fminf(by0, my); by1 = fmaxf(by1, my);
			}
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		bx0 = fminf(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
		by0 = fminf(by0, __shfl_xor_sync(0xffffffffu, by0, o));
		bx1 = fmaxf(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
		by1 = fmaxf(by1, __shfl_xor_sync(0xffffffffu, by1, o));
	}
	if (lane == 0 && bx0 <= bx1) {
		// fminf / fmaxf drop NaNs; huge or NaN-only boxes fail the size test below
		const bool sane = bx0 > -1.0e6f && bx1 < 1.0e6f && by0 > -1.0e6f && by1 < 1.0e6f;
		atomicMin(&s_box[0], sane ? (int)floorf(bx0) : INT_MIN);
		atomicMin(&s_box[1], sane ? (int)floorf(by0) : INT_MIN);
		atomicMax(&s_box[2], sane ? (int)floorf(bx1) : INT_MAX);
		atomicMax(&s_box[3], sane ? (int)floorf(by1) : INT_MAX);
	}
	__syncthreads();
	// window [wx0, wx0 + WP) x [wy0, wy0 + WH): taps ix .. ix+1 of every sample plus one pixel of
	// margin; wx0 is rounded down to 16 so that rows are copied as aligned 16-byte chunks
	const long long bw = (long long)s_box[2] - s_box[0], bh = (long long)s_box[3] - s_box[1];
	const bool box_ok = s_box[0] <= s_box[2] && bw < 4096 && bh < 4096;
	const int wx0 = box_ok ? ((s_box[0] - 1) & ~15) : 0, wy0 = box_ok ? s_box[1] - 1 : 0;
	const int WP = box_ok ? (((s_box[2] + 3 - wx0) + 15) & ~15) : 0, WH = box_ok ? s_box[3] + 3 - wy0 : 0;
	const bool staged = box_ok && WP * WH <= WIN_BYTES && ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)spitch) & 15) == 0;
	if (staged) {
		const int cpr = WP >> 4;  // 16-byte chunks per window row
		for (int c = tid; c < cpr * WH; c += TBX * TBY) {
			const int row = c / cpr, cx = (c - row * cpr) << 4;
			const int y = wy0 + row, x = wx0 + cx;
			// bytes of this chunk that lie inside the image; cp.async zero-fills the rest
			const int valid = ((unsigned)y < (unsigned)H && x >= 0) ? max(0, min(16, W - x)) : 0;
			og_cp_async16(s_win + row * WP + cx, valid > 0 ? src + y * spitch + x : src, valid);
		}
		asm volatile("cp.async.commit_group;" ::: "memory");
		asm volatile("cp.async.wait_group 0;" ::: "memory");
	}
	__syncthreads();
	if (!active) return;

	// ---- pass B: image
	uint8_t *orow = p.dst + (int64_t)ybeg * p.dst_pitch + xa;
	for (int yy = 0; yy < nrows; ++yy) {
		float2 d[TPX];
		eval_row(yy, d);
		float mx[TPX], my[TPX];
#pragma unroll
		for (int k = 0; k < TPX; ++k) {
			mx[k] = SCALED ? __fmul_rn(d[k].x, scx) : d[k].x;
			my[k] = SCALED ? __fmul_rn(d[k].y, scy) : d[k].y;
		}
		int v[TPX];
		int ox[TPX], oy[TPX];
		og_fix q[TPX];
		bool all_in = staged;
#pragma unroll
		for (int k = 0; k < TPX; ++k) {
			const int sxq = __float2int_rn(__fmul_rn(mx[k], 32.f));
			const int syq = __float2int_rn(__fmul_rn(my[k], 32.f));
			q[k].fx = sxq & 31;
			q[k].fy = syq & 31;
			ox[k] = (sxq >> 5) - wx0;
			oy[k] = (syq >> 5) - wy0;
			all_in = all_in && (unsigned)ox[k] < (unsigned)(WP - 1) && (unsigned)oy[k] < (unsigned)(WH - 1);
		}
		if (all_in) {
			int p00[TPX], p01[TPX], p10[TPX], p11[TPX];
#pragma unroll
			for (int k = 0; k < TPX; ++k) {
				const uint8_t *t = s_win + (oy[k] * WP + ox[k]);
				p00[k] = t[0];
				p01[k] = t[1];
				p10[k] = t[WP];
				p11[k] = t[WP + 1];
			}
#pragma unroll
			for (int k = 0; k < TPX; ++k) v[k] = og_blend(q[k], p00[k], p01[k], p10[k], p11[k]);
		} else {
			og_sample_linear_batch<TPX>(src, H, W, spitch, mx, my, v);
		}
		store_quad(orow, v, na, nb, img_al);
		orow += p.dst_pitch;
	}
}

// ---------------------------------------------------------------------------------------
// Pipelined variant (LINEAR image output): two pixels per thread instead of four, which halves
// the per-thread x-axis state and leaves room for a second tap buffer: the taps of row r are
// requested, then row r-1 is blended and stored while they are in flight (in og_page_kernel the
// blend waits on its own 16 loads: 41 % of that kernel's stall samples sit on the first multiply
// after them).  A warp covers 64 consecutive pixels of a row, lane l the pair {2l, 2l+1}: float4
// map stores are 512 contiguous bytes per warp, 16-bit image stores 64.
// ---------------------------------------------------------------------------------------
#ifndef OG_P2_ROWS
#define OG_P2_ROWS 16
#endif
#ifndef OG_P2_TBY
#define OG_P2_TBY 2
#endif
#ifndef OG_P2_MINB
#define OG_P2_MINB 4
#endif
#ifndef OG_P2_PREFETCH
#define OG_P2_PREFETCH 1
#endif
constexpr int P2_TBX = 128;            // threads along x: 256 pixels per CTA row
constexpr int P2_TBY = OG_P2_TBY;      // thread rows
constexpr int P2_ROWS = OG_P2_ROWS;    // rows walked per thread
constexpr int P2_TILE_W = 2 * P2_TBX;
constexpr int P2_TILE_H = P2_TBY * P2_ROWS;

template <bool SCALED, bool WRITE_MAP, bool TEX>
__global__ void __launch_bounds__(P2_TBX *P2_TBY, OG_P2_MINB) og_page_pipelined_kernel(const __grid_constant__ PageParams p) {
	__shared__ RowAxis s_rows[P2_TILE_H];
	__shared__ int s_win[4];
	const int Wd = p.gw * p.cell, Hd = p.gh * p.cell;
	const int ytile = blockIdx.y * P2_TILE_H;
	const double inv_cell = 1.0 / (double)p.cell;
	const int tid = threadIdx.y * P2_TBX + threadIdx.x;
	constexpr bool PREFETCH = OG_P2_PREFETCH && WRITE_MAP;
	constexpr int PF_WARP = (P2_TILE_H + 31) / 32;  // the first warp that does not fill s_rows
	static_assert(PF_WARP * 32 < P2_TBX * P2_TBY, "no warp left for the prefetch window");
	if (tid < P2_TILE_H) {
		og_axis ay = og_resize_axis(min(ytile + tid, Hd - 1), inv_cell);
		RowAxis ra;
		ra.r0 = min(max(ay.i, 0), p.gh - 1);
		ra.r1 = min(max(ay.i + 1, 0), p.gh - 1);
		ra.b1 = ay.f;
		ra.b0 = __fsub_rn(1.f, ay.f);
		s_rows[tid] = ra;
	} else if (PREFETCH && (tid >> 5) == PF_WARP) {
		// The source window of the tile (bounding box of nine map samples), for the prefetch below
		const int l = tid & 31, q = min(l, 8);
		const int px = min(blockIdx.x * P2_TILE_W + (q % 3) * (P2_TILE_W / 2), Wd - 1);
		const int py = min(ytile + (q / 3) * (P2_TILE_H / 2), Hd - 1);
		const float2 v = og_dense_node(p.grid, p.gh, p.gw, inv_cell, py, px);
		const float vx = SCALED ? __fmul_rn(v.x, p.scale_x) : v.x, vy = SCALED ? __fmul_rn(v.y, p.scale_y) : v.y;
		const bool ok = fabsf(vx) < 1.0e6f && fabsf(vy) < 1.0e6f;  // false for NaN
		int bx0 = ok ? (int)floorf(vx) - 2 : INT_MAX, by0 = ok ? (int)floorf(vy) - 2 : INT_MAX;
		int bx1 = ok ? (int)floorf(vx) + 3 : INT_MIN, by1 = ok ? (int)floorf(vy) + 3 : INT_MIN;
#pragma unroll
		for (int o = 8; o > 0; o >>= 1) {
			bx0 = min(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
			by0 = min(by0, __shfl_xor_sync(0xffffffffu, by0, o));
			bx1 = max(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
			by1 = max(by1, __shfl_xor_sync(0xffffffffu, by1, o));
		}
		if (l == 0) {
			s_win[0] = bx0; s_win[1] = by0; s_win[2] = bx1; s_win[3] = by1;
		}
	}
	// x-axis state (independent of the shared tables: ahead of the barrier)
	const int x0 = (blockIdx.x * P2_TBX + threadIdx.x) * 2;
	const int gw = p.gw;
	int sx[2], sx1[2];
	float a0[2], a1[2];
#pragma unroll
	for (int k = 0; k < 2; ++k) {
		og_axis ax = og_resize_axis_x(min(x0 + k, Wd - 1), inv_cell, gw);
		sx[k] = ax.i;
		sx1[k] = min(ax.i + 1, gw - 1);
		a1[k] = ax.f;
		a0[k] = __fsub_rn(1.f, ax.f);
	}
	__syncthreads();
	if (PREFETCH) {
		// The window is requested into L2 before the first tap: the 281 MB map stream keeps DRAM
		// busy with writes, and a row's taps that miss L2 wait behind them (every output row
		// touches a new source row).
		const int bx0 = max(s_win[0], 0), by0 = max(s_win[1], 0), bx1 = min(s_win[2], p.W - 1), by1 = min(s_win[3], p.H - 1);
		if (bx0 <= bx1 && by0 <= by1 && by1 - by0 < 4 * P2_TILE_H && bx1 - bx0 < 4 * P2_TILE_W) {
			const uintptr_t base = reinterpret_cast<uintptr_t>(p.src);
			for (int y = by0 + (tid >> 3); y <= by1; y += (P2_TBX * P2_TBY) >> 3) {
				const uintptr_t a = (base + (uintptr_t)((int64_t)y * p.src_pitch + bx0)) & ~(uintptr_t)127;
				const uintptr_t z = base + (uintptr_t)((int64_t)y * p.src_pitch + bx1);
				for (uintptr_t q = a + (tid & 7) * 128; q <= z; q += 8 * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
			}
		}
	}
	const int ybeg = ytile + threadIdx.y * P2_ROWS;
	if (x0 >= Wd || ybeg >= Hd) return;
	const float2 *__restrict__ grid = p.grid;
	const int npx = min(2, Wd - x0);
	const bool map_al = WRITE_MAP && npx == 2 && (Wd & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.map_out) & 15) == 0);
	const bool img_al = npx == 2 && (p.dst_pitch & 1) == 0 && ((reinterpret_cast<uintptr_t>(p.dst) & 1) == 0);
	const uint8_t *__restrict__ src = p.src;
	const int H = p.H, W = p.W, spitch = (int)p.src_pitch;
	const float scx = p.scale_x, scy = p.scale_y;

	int cur_r0 = -1, cur_r1 = -1;
	float2 h0[2], h1[2];
	float *mrow = WRITE_MAP ? p.map_out + ((size_t)ybeg * Wd + x0) * 2 : nullptr;
	uint8_t *orow = p.dst + (int64_t)ybeg * p.dst_pitch + x0;
	const RowAxis *rows = s_rows + threadIdx.y * P2_ROWS;
	const int nrows = min(P2_ROWS, Hd - ybeg);

	// the row loop, compiled once for the aligned two-pixel case (vector stores, no per-row
	// alignment branches) and once for the general one
	auto run = [&](auto aligned_tag) {
		constexpr bool AL = decltype(aligned_tag)::value;
		// map of one row (streamed out when wanted) and the request for its taps
		auto request = [&](int yy, og_taps (&t)[2]) {
			const RowAxis ra = rows[yy];
			if (ra.r0 != cur_r0 || ra.r1 != cur_r1) {  // next sparse cell: uniform across the CTA row
				cur_r0 = ra.r0;
				cur_r1 = ra.r1;
				const float2 *g0 = grid + ra.r0 * gw;
				const float2 *g1 = grid + ra.r1 * gw;
#pragma unroll
				for (int k = 0; k < 2; ++k) {
					const float2 s00 = __ldg(g0 + sx[k]), s01 = __ldg(g0 + sx1[k]);
					const float2 s10 = __ldg(g1 + sx[k]), s11 = __ldg(g1 + sx1[k]);
					h0[k].x = og_lerp_cv(s00.x, s01.x, a0[k], a1[k]);
					h0[k].y = og_lerp_cv(s00.y, s01.y, a0[k], a1[k]);
					h1[k].x = og_lerp_cv(s10.x, s11.x, a0[k], a1[k]);
					h1[k].y = og_lerp_cv(s10.y, s11.y, a0[k], a1[k]);
				}
			}
			float2 d[2];
#pragma unroll
			for (int k = 0; k < 2; ++k) {
				d[k].x = og_lerp_cv(h0[k].x, h1[k].x, ra.b0, ra.b1);
				d[k].y = og_lerp_cv(h0[k].y, h1[k].y, ra.b0, ra.b1);
			}
			if (WRITE_MAP) {
				if (AL) __stcs(reinterpret_cast<float4 *>(mrow), make_float4(d[0].x, d[0].y, d[1].x, d[1].y));
				else {
					__stcs(reinterpret_cast<float2 *>(mrow), d[0]);
					if (npx > 1) __stcs(reinterpret_cast<float2 *>(mrow) + 1, d[1]);
				}
				mrow += (size_t)Wd * 2;
			}
			float mx[2], my[2];
#pragma unroll
			for (int k = 0; k < 2; ++k) {
				mx[k] = SCALED ? __fmul_rn(d[k].x, scx) : d[k].x;
				my[k] = SCALED ? __fmul_rn(d[k].y, scy) : d[k].y;
			}
			if (TEX) {
				// the four taps of a sample as one gather (TLD4): texels (ix .. ix + 1, iy .. iy + 1) around
				// (ix + 1, iy + 1), zero outside the image = cv2.remap's BORDER_CONSTANT 0
#pragma unroll
				for (int k = 0; k < 2; ++k) {
					const int fsx = __float2int_rn(__fmul_rn(mx[k], 32.f)), fsy = __float2int_rn(__fmul_rn(my[k], 32.f));
					t[k].fx = fsx & 31;
					t[k].fy = fsy & 31;
					const float tx = __int2float_rn(og_sat_short(fsx >> 5) + 1), ty = __int2float_rn(og_sat_short(fsy >> 5) + 1);
					const uchar4 g = tex2Dgather<uchar4>(p.tex, tx, ty, 0);
					t[k].p00 = g.w;
					t[k].p01 = g.z;
					t[k].p10 = g.x;
					t[k].p11 = g.y;
				}
			} else {
				og_load_taps_batch<2>(src, H, W, spitch, mx, my, t);
			}
		};
		auto finish = [&](const og_taps (&t)[2]) {
			const int v0 = og_blend_taps(t[0]), v1 = og_blend_taps(t[1]);
			if (AL) __stcs(reinterpret_cast<unsigned short *>(orow), (unsigned short)(v0 | (v1 << 8)));
			else {
				orow[0] = (uint8_t)v0;
				if (npx > 1) orow[1] = (uint8_t)v1;
			}
			orow += p.dst_pitch;
		};

		og_taps ta[2], tb[2];
		request(0, ta);
		int yy = 1;
		for (; yy + 1 < nrows; yy += 2) {
			request(yy, tb);
			finish(ta);
			request(yy + 1, ta);
			finish(tb);
		}
		if (yy < nrows) {
			request(yy, tb);
			finish(ta);
			finish(tb);
		} else {
			finish(ta);
		}
	};
	if ((!WRITE_MAP || map_al) && img_al) run(std::true_type{});
	else run(std::false_type{});
}

template <bool LINEAR, bool SCALED>
void launch_variant(const PageParams &p, bool stage, dim3 grid, dim3 block, cudaStream_t st) {
	const int Wd = p.gw * p.cell, Hd = p.gh * p.cell;
	const dim3 block2(P2_TBX, P2_TBY), grid2((Wd + P2_TILE_W - 1) / P2_TILE_W, (Hd + P2_TILE_H - 1) / P2_TILE_H);
	const bool pipelined = LINEAR && !stage && OG_PAGE_PIPELINED;
	if (pipelined && p.tex != 0 && p.dst != nullptr) {
		if (p.map_out != nullptr) og_page_pipelined_kernel<SCALED, true, true><<<grid2, block2, 0, st>>>(p);
		else og_page_pipelined_kernel<SCALED, false, true><<<grid2, block2, 0, st>>>(p);
	} else if (p.map_out != nullptr && p.dst != nullptr) {
		if (pipelined) og_page_pipelined_kernel<SCALED, true, false><<<grid2, block2, 0, st>>>(p);
		else if (LINEAR && stage) og_page_staged_kernel<SCALED, true><<<grid, block, 0, st>>>(p);
		else og_page_kernel<LINEAR, SCALED, true, true><<<grid, block, 0, st>>>(p);
	} else if (p.dst != nullptr) {
		if (pipelined) og_page_pipelined_kernel<SCALED, false, false><<<grid2, block2, 0, st>>>(p);
		else if (LINEAR && stage) og_page_staged_kernel<SCALED, false><<<grid, block, 0, st>>>(p);
		else og_page_kernel<LINEAR, SCALED, false, true><<<grid, block, 0, st>>>(p);
	} else {
		og_page_kernel<true, false, true, false><<<grid, block, 0, st>>>(p);
	}
}

int launch_page(const PageParams &p, int interp, cudaStream_t st) {
	const int Wd = p.gw * p.cell, Hd = p.gh * p.cell;
	dim3 block(TBX, TBY);
	dim3 grid((Wd + TILE_W - 1) / TILE_W, (Hd + TILE_H - 1) / TILE_H);
	const bool scaled = !(p.scale_x == 1.f && p.scale_y == 1.f);
	const bool stage = (interp & OG_FLAG_STAGE_SMEM) != 0 || OG_PAGE_STAGE;
	if ((interp & 0xff) == OG_INTER_LINEAR) {
		if (scaled) launch_variant<true, true>(p, stage, grid, block, st);
		else launch_variant<true, false>(p, stage, grid, block, st);
	} else {
		if (scaled) launch_variant<false, true>(p, stage, grid, block, st);
		else launch_variant<false, false>(p, stage, grid, block, st);
	}
	return og_check_launch("og_page_kernel");
}

// OG_PAGE_TEX experiment: texture objects over the caller's source images, kept per (device, pointer,
// shape, pitch) -- a descriptor only holds the address, so it stays valid when the memory is rewritten
// or reallocated at the same place; nothing is destroyed while kernels may still use it.
cudaTextureObject_t page_texture(const uint8_t *src, int H, int W, int64_t pitch) {
	static const bool enabled = getenv("OG_PAGE_TEX") != nullptr;
	if (!enabled || (reinterpret_cast<uintptr_t>(src) & 511) != 0 || (pitch & 31) != 0) return 0;
	struct Entry { int dev; const uint8_t *src; int H, W; int64_t pitch; cudaTextureObject_t tex; };
	static Entry cache[64];
	static int used = 0;
	static std::mutex mu;
	std::lock_guard<std::mutex> lock(mu);
	const int dev = og_device_ordinal();
	for (int i = 0; i < used; ++i)
		if (cache[i].dev == dev && cache[i].src == src && cache[i].H == H && cache[i].W == W && cache[i].pitch == pitch) return cache[i].tex;
	if (used == 64) return 0;
	cudaResourceDesc rd{};
	rd.resType = cudaResourceTypePitch2D;
	rd.res.pitch2D.devPtr = const_cast<uint8_t *>(src);
	rd.res.pitch2D.desc = cudaCreateChannelDesc<unsigned char>();
	rd.res.pitch2D.width = (size_t)W;
	rd.res.pitch2D.height = (size_t)H;
	rd.res.pitch2D.pitchInBytes = (size_t)pitch;
	cudaTextureDesc td{};
	td.addressMode[0] = td.addressMode[1] = cudaAddressModeBorder;
	td.filterMode = cudaFilterModePoint;
	td.readMode = cudaReadModeElementType;
	td.normalizedCoords = 0;
	cudaTextureObject_t tex = 0;
	if (cudaCreateTextureObject(&tex, &rd, &td, nullptr) != cudaSuccess) {
		cudaGetLastError();
		return 0;
	}
	cache[used++] = Entry{dev, src, H, W, pitch, tex};
	return tex;
}

int check_grid(const float *grid, int gh, int gw, int cell) {
	OG_REQUIRE(grid != nullptr, OG_E_ARG, "grid is NULL");
	OG_REQUIRE(gh >= 2 && gw >= 2 && cell >= 1, OG_E_ARG, "grid shape (%d, %d) cell %d invalid", gh, gw, cell);
	OG_REQUIRE((reinterpret_cast<uintptr_t>(grid) & 7) == 0, OG_E_ALIGN, "grid must be 8-byte aligned");
	// cv2.remap asserts dst.cols < SHRT_MAX && dst.rows < SHRT_MAX
	OG_REQUIRE((int64_t)gh * cell < 32767 && (int64_t)gw * cell < 32767, OG_E_RANGE,
		"dewarped size %lld x %lld exceeds OpenCV's 32767 limit", (long long)gw * cell, (long long)gh * cell);
	return OG_OK;
}

}  // namespace

extern "C" int og_expand_map(const float *grid, int gh, int gw, int cell, float *map_out, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	OG_REQUIRE(map_out != nullptr, OG_E_ARG, "map_out is NULL");
	OG_REQUIRE((reinterpret_cast<uintptr_t>(map_out) & 7) == 0, OG_E_ALIGN, "map_out must be 8-byte aligned");
	PageParams p{};
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh;
	p.gw = gw;
	p.cell = cell;
	p.map_out = map_out;
	p.scale_x = p.scale_y = 1.f;
	return launch_page(p, OG_INTER_LINEAR, (cudaStream_t)stream);
}

extern "C" int og_remap_page(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell,
	uint8_t *dst, int64_t dst_pitch, float *map_out,
	float scale_x, float scale_y, int interp, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	OG_REQUIRE(dst != nullptr || map_out != nullptr, OG_E_ARG, "both dst and map_out are NULL");
	OG_REQUIRE((interp & 0xff) == OG_INTER_LINEAR || (interp & 0xff) == OG_INTER_NEAREST, OG_E_ARG, "interp %d unsupported", interp);
	OG_REQUIRE((interp & ~(0xff | OG_FLAG_STAGE_SMEM)) == 0, OG_E_ARG, "interp %d has unknown flag bits", interp);
	if (dst != nullptr) {
		OG_REQUIRE(src != nullptr && H > 0 && W > 0, OG_E_ARG, "src is NULL or empty (%d x %d)", W, H);
		OG_REQUIRE(H < 32767 && W < 32767, OG_E_RANGE, "source %d x %d exceeds OpenCV's 32767 limit", W, H);
		OG_REQUIRE(src_pitch >= W && dst_pitch >= (int64_t)gw * cell, OG_E_ARG, "pitch smaller than row");
		OG_REQUIRE((int64_t)H * src_pitch < ((int64_t)1 << 31), OG_E_RANGE, "source image larger than 2 GiB");
	}
	if (map_out != nullptr)
		OG_REQUIRE((reinterpret_cast<uintptr_t>(map_out) & 7) == 0, OG_E_ALIGN, "map_out must be 8-byte aligned");
	PageParams p{};
	p.src = src;
	p.H = H;
	p.W = W;
	p.src_pitch = src_pitch;
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh;
	p.gw = gw;
	p.cell = cell;
	p.dst = dst;
	p.dst_pitch = dst_pitch;
	p.map_out = map_out;
	p.scale_x = scale_x;
	p.scale_y = scale_y;
	p.tex = (dst != nullptr && (interp & 0xff) == OG_INTER_LINEAR) ? page_texture(src, H, W, src_pitch) : 0;
	return launch_page(p, interp, (cudaStream_t)stream);
}
