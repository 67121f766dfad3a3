// Warp-autonomous Sauvola for windows up to 15 (round 2): no CTA barrier, every source byte read
// once per stripe, window sums from byte dot products.
//
// skimage 0.16.2 threshold_sauvola (origami/core/binarize.py:26-33, page.py:101-104,
// batch/detect/layout.py:157-169): reflect-padded centred w x w window, m = S1 / w^2,
// s = sqrt(max(S2 / w^2 - m m, 0)), t = m (1 + k (s / r - 1)) in float64, pixel > t.  S1 and S2
// (sums of p and p^2) are exact integers here, the comparison is screened in float32 and the few
// pixels within 1e-3 gray levels of their threshold take the reference's float64 sequence --
// the same decision procedure as sauvola_tile in og_binarize.cu, only the sums are formed differently:
//
//   * a warp owns a stripe of columns and a run of rows; lane l holds PX = 4 or 8 neighbouring pixels
//     of each row as one or two 32-bit words (one coalesced 128- or 256-byte row read per warp); the
//     lanes whose windows reach past the warp's columns (8 pixels on either side) produce no output,
//     so a stripe is 112 (PX = 4) or 240 (PX = 8) output columns wide;
//   * horizontal: the neighbouring lanes' words come by shuffle (raw pixels, four shuffles), and the
//     window sums of p and p^2 of the lane's outputs are dp4a dot products of those words with byte
//     masks (p: word . 0x01010101, p^2: (word & mask) . word), the words common to all windows
//     (and to the windows of a group of four outputs) summed once -- six to seven instructions per
//     pixel for both sums;
//   * vertical: the four (S1, S2) pairs of the lane run down the rows, + the row entering the window,
//     - the row that left it.  A row's horizontal sums wait for that in a shared-memory ring of
//     `window` rows (zero at the start of a run), packed as H2 << 12 | H1 (H1 <= 15 * 255 < 2^12, H2 <= 15 * 255^2 < 2^20:
//     exactly 32 bits, which is what limits this kernel to windows <= 15), one 16-byte store and
//     load per lane and row.  The ring is private to the warp: no barrier anywhere.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

constexpr int WS_WARPS = 4;  // warps per CTA (independent of each other)
#ifndef OG_WS_AHEAD
#define OG_WS_AHEAD 1
#endif
#ifndef OG_WS_AHEAD_STRIPS
#define OG_WS_AHEAD_STRIPS 2
#endif
constexpr int WS_AHEAD = OG_WS_AHEAD;  // rows requested ahead of use

template <int PX>
struct WsGeom {
	static constexpr int NW = PX / 4;              // words per lane and row
	static constexpr int NK = NW + 4;              // words a lane's windows can touch: k = -2 .. NW + 1
	static constexpr int EDGE = 8 / PX;            // lanes at either end of the warp without output
	static constexpr int OUT = (32 - 2 * EDGE) * PX;  // output columns per warp
};

__device__ __noinline__ int ws_reflect_far(int i, int n) {  // numpy.pad(mode="reflect"), any distance
	if (n == 1) return 0;
	const int period = 2 * (n - 1);
	i %= period;
	if (i < 0) i += period;
	return i >= n ? period - i : i;
}

__device__ __forceinline__ int ws_reflect(int i, int n) {
	if ((unsigned)i < (unsigned)n) return i;
	const int j = i < 0 ? -i : 2 * (n - 1) - i;
	if ((unsigned)j < (unsigned)n) return j;
	return ws_reflect_far(i, n);
}

// bytes of word k (columns 4 k .. 4 k + 3 relative to the lane's first pixel, k = -2 .. NW + 1) that
// lie in the window of the lane's output j
template <int RAD>
__device__ __forceinline__ constexpr unsigned ws_mask(int j, int k) {
	unsigned m = 0;
	for (int b = 0; b < 4; ++b) {
		const int d = 4 * k + b - j;
		if (d >= -RAD && d <= RAD) m |= 1u << (8 * b);
	}
	return m;
}

template <int RAD, int PX>
__device__ __forceinline__ constexpr unsigned ws_all(int k) {  // ... in the windows of all outputs
	unsigned m = 0x01010101u;
	for (int j = 0; j < PX; ++j) m &= ws_mask<RAD>(j, k);
	return m;
}

template <int RAD, int PX>
__device__ __forceinline__ constexpr unsigned ws_group(int g, int k) {  // ... of outputs 4 g .. 4 g + 3, beyond ws_all
	unsigned m = 0x01010101u;
	for (int j = 4 * g; j < 4 * g + 4; ++j) m &= ws_mask<RAD>(j, k);
	return m & ~ws_all<RAD, PX>(k);
}

__device__ __forceinline__ void ws_dot(unsigned word, unsigned mask, unsigned &s1, unsigned &s2) {
	s1 = __dp4a(word, mask, s1);
	s2 = __dp4a(mask == 0x01010101u ? word : (word & (mask * 255u)), word, s2);
}

// window sums of p (h1) and p^2 (h2) along the row for the lane's PX outputs; wd[k + 2] = word k
template <int RAD, int PX>
__device__ __forceinline__ void ws_hsums(const unsigned (&wd)[WsGeom<PX>::NK], unsigned (&h1)[PX], unsigned (&h2)[PX]) {
	constexpr int NK = WsGeom<PX>::NK;
	unsigned b1 = 0, b2 = 0;
#pragma unroll
	for (int k = 0; k < NK; ++k)
		if (ws_all<RAD, PX>(k - 2) != 0) ws_dot(wd[k], ws_all<RAD, PX>(k - 2), b1, b2);
#pragma unroll
	for (int g = 0; g < PX / 4; ++g) {
		unsigned g1 = b1, g2 = b2;
#pragma unroll
		for (int k = 0; k < NK; ++k)
			if (ws_group<RAD, PX>(g, k - 2) != 0) ws_dot(wd[k], ws_group<RAD, PX>(g, k - 2), g1, g2);
#pragma unroll
		for (int j = 4 * g; j < 4 * g + 4; ++j) {
			unsigned e1 = g1, e2 = g2;
#pragma unroll
			for (int k = 0; k < NK; ++k) {
				const unsigned m = ws_mask<RAD>(j, k - 2) & ~ws_all<RAD, PX>(k - 2) & ~ws_group<RAD, PX>(g, k - 2);
				if (m != 0) ws_dot(wd[k], m, e1, e2);
			}
			h1[j] = e1;
			h2[j] = e2;
		}
	}
}

// prmt.b32, generic mode: a selector nibble with its top bit set replicates the sign bit of the
// selected byte over the target byte (__byte_perm masks that bit away)
__device__ __forceinline__ unsigned ws_prmt(unsigned a, unsigned b, unsigned sel) {
	unsigned d;
	asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
	return d;
}

struct WsConst {
	float fc1, fc2;  // (1 - k) / n and k / (n^2 r) of the float32 screen
	double n, kk, rr;
};

// the pixels of a group of four whose float32 estimate is within 1e-3 of the pixel: the reference's
// float64 sequence decides them
__device__ __noinline__ unsigned ws_settle(unsigned res, unsigned cw, uint4 s1, uint4 s2, unsigned undecided, double n, double kk, double rr) {
	const unsigned a1[4] = {s1.x, s1.y, s1.z, s1.w}, a2[4] = {s2.x, s2.y, s2.z, s2.w};
#pragma unroll
	for (int j = 0; j < 4; ++j) {
		if (!(undecided >> j & 1u)) continue;
		const double m = __ddiv_rn((double)a1[j], n);
		const double g2 = __ddiv_rn((double)a2[j], n);
		double var = __dsub_rn(g2, __dmul_rn(m, m));
		var = var > 0.0 ? var : 0.0;
		const double sd = __dsqrt_rn(var);
		const double t = __dmul_rn(m, __dadd_rn(1.0, __dmul_rn(kk, __dsub_rn(__ddiv_rn(sd, rr), 1.0))));
		const bool white = (double)((cw >> (8 * j)) & 255u) > t;
		res = (res & ~(255u << (8 * j))) | (white ? 255u << (8 * j) : 0u);
	}
	return res;
}

// One stripe: output columns x0 .. x0 + OUT - 1, output rows y0 .. y1 - 1 of an (h, w) image
// (pitch and out_pitch below 2^31).  ring_h: WIN x NW x 32 uint4, ring_w: (RAD + 1) x NW x 32 words.
template <int RAD, int PX, bool ALIGNED, int AHEAD = WS_AHEAD>
__device__ __forceinline__ void ws_stripe(const uint8_t *__restrict__ img, int h, int w, unsigned pitch,
	uint8_t *__restrict__ out, unsigned out_pitch, int x0, int y0, int y1, const WsConst c,
	uint4 *__restrict__ ring_h, unsigned *__restrict__ ring_w) {
	static_assert(ALIGNED || PX == 4, "unaligned rows are read one word per lane");
	using G = WsGeom<PX>;
	constexpr int WIN = 2 * RAD + 1, NW = G::NW, NK = G::NK, RW = RAD + 1;
	constexpr unsigned NWIN = WIN * WIN;
	const int lane = threadIdx.x & 31;
	const int cx = x0 - 8 + PX * lane;
	const bool inside = cx >= 0 && cx + PX - 1 < w;
	const bool writes = lane >= G::EDGE && lane < 32 - G::EDGE && cx < w;
	const bool write_words = ALIGNED && writes && cx + PX - 1 < w;
	const uint8_t *img_lane = img + (inside ? cx : 0);
	uint8_t *out_lane = out + cx;
	uint4 *rh = ring_h + lane;
	unsigned *rw = ring_w + lane;
#pragma unroll
	for (int i = 0; i < WIN * NW; ++i) rh[i * 32] = make_uint4(0u, 0u, 0u, 0u);  // the rows "before" the first leave as zeros
	// A lane of a stripe at the left or right border of the image holds reflected columns.  Four
	// neighbouring columns reflect to at most four neighbouring columns, so they lie in two aligned words:
	// the lane requests those and permutes them into place when the row is used, so that its loads
	// wait in the queue like everybody's.  Unaligned images (strips: a row starts wherever the previous
	// one ended) do the same in every lane: two aligned words and a selector that depends on the row.
	struct Row { unsigned w[NW], x[NW], sel; };
	unsigned boff[NW], bsel[NW];  // lowest column (aligned images: rounded down to a word) and byte selectors
	unsigned umax = 3;            // unaligned: the largest selector nibble before the row's shift
#pragma unroll
	for (int q = 0; q < NW; ++q) {
		boff[q] = inside ? (unsigned)(cx + 4 * q) : 0u;
		bsel[q] = 0x3210u;
		if (!inside) {
			int col[4], lo = 0x7fffffff, hi = 0;
#pragma unroll
			for (int b = 0; b < 4; ++b) {
				col[b] = ws_reflect(cx + 4 * q + b, w);
				lo = min(lo, col[b]);
				hi = max(hi, col[b]);
			}
			boff[q] = ALIGNED ? ((unsigned)lo & ~3u) : (unsigned)lo;
			bsel[q] = (unsigned)(col[0] - (int)boff[q]) | (unsigned)(col[1] - (int)boff[q]) << 4 |
				(unsigned)(col[2] - (int)boff[q]) << 8 | (unsigned)(col[3] - (int)boff[q]) << 12 |
				(hi >= (int)boff[q] + 4 ? 0x10000u : 0u);  // bit 16: the second word is needed (and holds a column of the image)
			if (!ALIGNED) umax = (unsigned)(hi - lo);
		}
	}
	auto load_row = [&](int r) -> Row {
		const int ry = (unsigned)r < (unsigned)h ? r : ws_reflect(r, h);
		Row v;
		v.sel = 0;
		if (!ALIGNED) {
			const uintptr_t a = reinterpret_cast<uintptr_t>(img + (uint64_t)(unsigned)ry * pitch + boff[0]);
			const unsigned *wp = reinterpret_cast<const unsigned *>(a & ~(uintptr_t)3);
			const unsigned sh = (unsigned)(a & 3);
			v.w[0] = __ldg(wp);
			v.x[0] = sh + umax >= 4 ? __ldg(wp + 1) : 0u;  // an aligned word that holds a byte of the image stays inside its page
			v.sel = (bsel[0] & 0xFFFFu) + sh * 0x1111u;
			return v;
		}
		const uint8_t *row = img_lane + (uint64_t)(unsigned)ry * pitch;
		if (inside) {
			if (PX == 8) {
				const uint2 t = __ldg(reinterpret_cast<const uint2 *>(row));
				v.w[0] = t.x;
				v.w[NW - 1] = t.y;
			} else {
				v.w[0] = __ldg(reinterpret_cast<const unsigned *>(row));
			}
		} else {
#pragma unroll
			for (int q = 0; q < NW; ++q) {
				const unsigned *wp = reinterpret_cast<const unsigned *>(row + boff[q]);
				v.w[q] = __ldg(wp);
				v.x[q] = (bsel[q] & 0x10000u) ? __ldg(wp + 1) : 0u;
			}
		}
		return v;
	};
	unsigned v1[PX], v2[PX];
#pragma unroll
	for (int j = 0; j < PX; ++j) v1[j] = v2[j] = 0;
	int slot = 0, wslot = 0;  // ring slots of row r (sums: WIN slots, pixels: RAD + 1 slots)
	auto body = [&](int r, Row cur, bool screen) {
		if (!ALIGNED) {
			cur.w[0] = __byte_perm(cur.w[0], cur.x[0], cur.sel);
		} else if (!inside) {
#pragma unroll
			for (int q = 0; q < NW; ++q) cur.w[q] = __byte_perm(cur.w[q], cur.x[q], bsel[q] & 0xFFFFu);
		}
		// the row entering the window replaces the row that left it WIN rows ago in the ring
		uint4 old[NW];
#pragma unroll
		for (int q = 0; q < NW; ++q) old[q] = rh[(slot * NW + q) * 32];
		unsigned wd[NK];
		if (PX == 4) {
			wd[2] = cur.w[0];
			wd[0] = __shfl_up_sync(0xffffffffu, wd[2], 2);
			wd[1] = __shfl_up_sync(0xffffffffu, wd[2], 1);
			wd[3] = __shfl_down_sync(0xffffffffu, wd[2], 1);
			wd[4] = __shfl_down_sync(0xffffffffu, wd[2], 2);
		} else {
			wd[2] = cur.w[0];
			wd[3] = cur.w[NW - 1];
			wd[0] = __shfl_up_sync(0xffffffffu, wd[2], 1);
			wd[1] = __shfl_up_sync(0xffffffffu, wd[3], 1);
			wd[NK - 2] = __shfl_down_sync(0xffffffffu, wd[2], 1);
			wd[NK - 1] = __shfl_down_sync(0xffffffffu, wd[3], 1);
		}
		unsigned h1[PX], h2[PX];
		ws_hsums<RAD, PX>(wd, h1, h2);
#pragma unroll
		for (int q = 0; q < NW; ++q) {
			const unsigned o[4] = {old[q].x, old[q].y, old[q].z, old[q].w};
#pragma unroll
			for (int j = 0; j < 4; ++j) {
				v1[4 * q + j] += h1[4 * q + j] - (o[j] & 0xFFFu);
				v2[4 * q + j] += h2[4 * q + j] - (o[j] >> 12);
			}
			rh[(slot * NW + q) * 32] = make_uint4(h2[4 * q] * 4096u + h1[4 * q], h2[4 * q + 1] * 4096u + h1[4 * q + 1],
				h2[4 * q + 2] * 4096u + h1[4 * q + 2], h2[4 * q + 3] * 4096u + h1[4 * q + 3]);
			rw[(wslot * NW + q) * 32] = cur.w[q];
		}
		slot = slot + 1 == WIN ? 0 : slot + 1;
		wslot = wslot + 1 == RW ? 0 : wslot + 1;  // now the slot of row r - RAD
		if (!screen) return;
		const int yo = r - RAD;
		const bool live = (unsigned)(yo - y0) < (unsigned)(y1 - y0);
		unsigned res[NW];
#pragma unroll
		for (int q = 0; q < NW; ++q) {
			const unsigned cw = rw[(wslot * NW + q) * 32];
			float diff[4];
#pragma unroll
			for (int j = 0; j < 4; ++j) {
				// D = n S2 - S1^2 exactly (n^2 255^2 < 2^32 for windows <= 15); the float32 estimate of the
				// threshold is within 2e-4 gray levels of the float64 one (see sauvola_tile), and the
				// approximate square root adds 6e-5 at most: |pixel - estimate| > 1e-3 decides
				const unsigned s1 = v1[4 * q + j];
				const unsigned D = NWIN * v2[4 * q + j] - s1 * s1;
				float rt;
				asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rt) : "f"((float)D));  // D is 0 or >= 1: nothing to flush
				const float fpv = __fsub_rn(__int_as_float((int)__byte_perm(cw, 0x4B000000u, 0x7650 + j)), 8388608.0f);
				diff[j] = __fmaf_rn(-(float)s1, __fmaf_rn(c.fc2, rt, c.fc1), fpv);
			}
			// bytes of sign bits: 0xFF where diff < 0 (or -0, which ws_settle decides)
			const unsigned n01 = ws_prmt(__float_as_uint(diff[0]), __float_as_uint(diff[1]), 0x00FBu);
			const unsigned n23 = ws_prmt(__float_as_uint(diff[2]), __float_as_uint(diff[3]), 0x00FBu);
			res[q] = ~__byte_perm(n01, n23, 0x5410);
			const float closest = fminf(fminf(fabsf(diff[0]), fabsf(diff[1])), fminf(fabsf(diff[2]), fabsf(diff[3])));
			if (!(closest > 1.0e-3f)) {
				unsigned undecided = 0;
#pragma unroll
				for (int j = 0; j < 4; ++j)
					if (!(fabsf(diff[j]) > 1.0e-3f)) undecided |= 1u << j;
				res[q] = ws_settle(res[q], cw, make_uint4(v1[4 * q], v1[4 * q + 1], v1[4 * q + 2], v1[4 * q + 3]),
					make_uint4(v2[4 * q], v2[4 * q + 1], v2[4 * q + 2], v2[4 * q + 3]), undecided, c.n, c.kk, c.rr);
			}
		}
		uint8_t *o = out_lane + (uint64_t)(unsigned)yo * out_pitch;
		if (write_words) {
			if (live) {
				if (PX == 8) __stcs(reinterpret_cast<uint2 *>(o), make_uint2(res[0], res[NW - 1]));
				else __stcs(reinterpret_cast<unsigned *>(o), res[0]);
			}
		} else if (writes && live) {
#pragma unroll
			for (int j = 0; j < PX; ++j)
				if (cx + j < w) o[j] = (uint8_t)(res[j >> 2] >> (8 * (j & 3)));
		}
	};
	// Rows are requested WS_AHEAD iterations before they are used (otherwise a warp has one row read in
	// flight and nothing to do until it lands); the loops are unrolled by that distance so that the
	// queue lives in fixed registers.  Rows past the end are harmless reflections whose outputs are
	// not stored; the first rows of a run only fill the window.
	const int r0 = y0 - RAD, r_end = y1 + RAD;
	constexpr int WARM = (2 * RAD) / AHEAD * AHEAD;
	Row ahead[AHEAD];
#pragma unroll
	for (int i = 0; i < AHEAD; ++i) ahead[i] = load_row(r0 + i);
	int r = r0;
	for (; r < r0 + WARM; r += AHEAD) {
#pragma unroll
		for (int u = 0; u < AHEAD; ++u) {
			const Row cur = ahead[u];
			ahead[u] = load_row(r + u + AHEAD);
			body(r + u, cur, false);
		}
	}
	for (; r < r_end; r += AHEAD) {
#pragma unroll
		for (int u = 0; u < AHEAD; ++u) {
			const Row cur = ahead[u];
			ahead[u] = load_row(r + u + AHEAD);
			body(r + u, cur, true);
		}
	}
}

template <int RAD, int PX>
__host__ __device__ constexpr size_t ws_warp_smem() {
	return (size_t)32 * (PX / 4) * ((2 * RAD + 1) * sizeof(uint4) + (RAD + 1) * sizeof(unsigned));
}

template <int RAD, int PX, int WARPS>
__device__ __forceinline__ void ws_rings(unsigned char *smem, uint4 *&ring_h, unsigned *&ring_w) {
	constexpr int NW = PX / 4, WIN = 2 * RAD + 1;
	const int wid = threadIdx.x >> 5;
	ring_h = reinterpret_cast<uint4 *>(smem) + wid * WIN * NW * 32;
	ring_w = reinterpret_cast<unsigned *>(smem + (size_t)WARPS * WIN * NW * 32 * sizeof(uint4)) + wid * (RAD + 1) * NW * 32;
}

// pages: item = (run of rows, stripe); neighbouring warps take neighbouring stripes of one run
template <int RAD, int PX, bool ALIGNED, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) og_sauvola_warp_page_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t src_pitch, uint8_t *__restrict__ dst, int64_t dst_pitch,
	const WsConst c, int seg_rows, int stripes, int items) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	uint4 *ring_h;
	unsigned *ring_w;
	ws_rings<RAD, PX, WARPS>(smem_raw, ring_h, ring_w);
	const int item = blockIdx.x * WARPS + (threadIdx.x >> 5);
	if (item >= items) return;
	const int seg = item / stripes, stripe = item - seg * stripes;
	const int y0 = seg * seg_rows;
	ws_stripe<RAD, PX, ALIGNED>(src, h, w, (unsigned)src_pitch, dst, (unsigned)dst_pitch, stripe * WsGeom<PX>::OUT, y0,
		min(h, y0 + seg_rows), c, ring_h, ring_w);
}

// strips: item = a 64-column chunk of the line plan; chunk q of a line works on stripe q when the
// line has that many stripes (a strip of w columns has ceil(w / 64) chunks >= ceil(w / 112) stripes).
// Items are handed out through the plan's work-ticket word (as in og_strips_screened_kernel): strips
// are short runs of unequal width, and a static deal leaves most warps idle behind the slowest
// (13 % of the warp slots busy, measured).  Every warp draws exactly one ticket past the end; the
// warp that draws the very last one puts the word back to zero.
template <int RAD>
__global__ void __launch_bounds__(WS_WARPS * 32) og_sauvola_warp_strips_kernel(
	const uint8_t *__restrict__ gray, const int32_t *__restrict__ steps, const int64_t *__restrict__ pix_offsets,
	const int32_t *__restrict__ chunk_offsets, int n, uint8_t *__restrict__ out, int64_t capacity, const WsConst c,
	int *__restrict__ ticket) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	uint4 *ring_h;
	unsigned *ring_w;
	ws_rings<RAD, 4, WS_WARPS>(smem_raw, ring_h, ring_w);
	const int total = __ldg(chunk_offsets + n);
	const int nwarps = gridDim.x * WS_WARPS;
	const int lane = threadIdx.x & 31;
	auto draw = [&]() {
		int t = 0;
		if (lane == 0) t = atomicAdd(ticket, 1);
		return __shfl_sync(0xffffffffu, t, 0);
	};
	int item = draw();
	while (item < total) {
		const int next = draw();  // the atomic's latency hides behind the item
		int lo = 0, hi = n;
		while (hi - lo > 1) {
			const int mid = (lo + hi) >> 1;
			if (__ldg(chunk_offsets + mid) <= item) lo = mid; else hi = mid;
		}
		const int q = item - __ldg(chunk_offsets + lo);
		const int w = __ldg(steps + 2 * lo), h = __ldg(steps + 2 * lo + 1);
		const int64_t off = __ldg(pix_offsets + lo);
		if (q * WsGeom<4>::OUT < w && h > 0 && off + (int64_t)w * h <= capacity) {
			__syncwarp();
			ws_stripe<RAD, 4, false, OG_WS_AHEAD_STRIPS>(gray + off, h, w, (unsigned)w, out + off, (unsigned)w, q * WsGeom<4>::OUT, 0, h, c, ring_h, ring_w);
			__syncwarp();
		}
		item = next;
	}
	if (lane == 0 && item == total + nwarps - 1) *ticket = 0;
}

WsConst ws_const(int window, double k, double r) {
	WsConst c;
	const double n = (double)window * window;
	c.fc1 = (float)((1.0 - k) / n);
	c.fc2 = (float)(k / (n * n * r));
	c.n = n;
	c.kk = k;
	c.rr = r;
	return c;
}

int ws_env_int(const char *name, int fallback) {
	const char *v = getenv(name);
	return v != nullptr && atoi(v) > 0 ? atoi(v) : fallback;
}

template <int RAD, int PX, bool ALIGNED, int WARPS>
int ws_launch_page_as(const uint8_t *src, int h, int w, int64_t src_pitch, uint8_t *dst, int64_t dst_pitch,
	double k, double r, int slot, cudaStream_t st) {
	constexpr size_t smem = WARPS * ws_warp_smem<RAD, PX>();
	const int stripes = (w + WsGeom<PX>::OUT - 1) / WsGeom<PX>::OUT;
	// One wave of runs: as many warps as fit beside each other (shared memory: the rings), each with a
	// run as long as that allows -- a run re-reads 2 RAD rows to fill its window.
	const int fit = (int)((size_t)220 * 1024 / (smem + 1024)) * WARPS;
	static const int per_sm_env = ws_env_int("OG_SAUVOLA_WARPS", 0);
	const int per_sm = per_sm_env > 0 ? per_sm_env : (fit < 48 ? fit : 48);
	int segments = max(1, og_num_sms() * per_sm / stripes);
	segments = max(1, min(segments, (h + 4 * (2 * RAD + 1) - 1) / (4 * (2 * RAD + 1))));
	const int seg_rows = (h + segments - 1) / segments;
	segments = (h + seg_rows - 1) / seg_rows;
	const int64_t items64 = (int64_t)stripes * segments;
	OG_REQUIRE(items64 < ((int64_t)1 << 30) && src_pitch < ((int64_t)1 << 31) && dst_pitch < ((int64_t)1 << 31), OG_E_RANGE, "image too large");
	const int items = (int)items64;
	auto kernel = og_sauvola_warp_page_kernel<RAD, PX, ALIGNED, WARPS>;
	if (og_first_use(slot))
		cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	kernel<<<(items + WARPS - 1) / WARPS, WARPS * 32, smem, st>>>(
		src, h, w, src_pitch, dst, dst_pitch, ws_const(2 * RAD + 1, k, r), seg_rows, stripes, items);
	return og_check_launch("og_sauvola_warp_page_kernel");
}

template <int RAD>
int ws_launch_page(const uint8_t *src, int h, int w, int64_t src_pitch, uint8_t *dst, int64_t dst_pitch,
	double k, double r, cudaStream_t st) {
	const uintptr_t bits = reinterpret_cast<uintptr_t>(src) | (uintptr_t)src_pitch | reinterpret_cast<uintptr_t>(dst) | (uintptr_t)dst_pitch;
	static const int px_env = ws_env_int("OG_SAUVOLA_PX", 8);
	if ((bits & 7) == 0 && px_env == 8 && w >= 256)
		return ws_launch_page_as<RAD, 8, true, 2>(src, h, w, src_pitch, dst, dst_pitch, k, r, 3 * RAD, st);
	if ((bits & 3) == 0)
		return ws_launch_page_as<RAD, 4, true, WS_WARPS>(src, h, w, src_pitch, dst, dst_pitch, k, r, 3 * RAD + 1, st);
	return ws_launch_page_as<RAD, 4, false, WS_WARPS>(src, h, w, src_pitch, dst, dst_pitch, k, r, 3 * RAD + 2, st);
}

template <int RAD>
int ws_launch_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, uint8_t *out, int64_t capacity, cudaStream_t st) {
	constexpr size_t smem = WS_WARPS * ws_warp_smem<RAD, 4>();
	if (og_first_use(48 + RAD))
		cudaFuncSetAttribute(og_sauvola_warp_strips_kernel<RAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	// the word behind the prefix sums is the plan's work ticket (zero between calls, see og_line_steps)
	int *ticket = const_cast<int *>(chunk_offsets) + n + 1;
	const int fit = (int)((size_t)220 * 1024 / (smem + 1024));
	og_sauvola_warp_strips_kernel<RAD><<<og_num_sms() * fit, WS_WARPS * 32, smem, st>>>(gray, steps, pix_offsets, chunk_offsets, n, out, capacity,
		ws_const(2 * RAD + 1, 0.2, 127.5), ticket);
	return og_check_launch("og_sauvola_warp_strips_kernel");
}

}  // namespace

#define OG_WS_SWITCH(CALL) \
	switch (window) { \
		case 3: return CALL(1); case 5: return CALL(2); case 7: return CALL(3); case 9: return CALL(4); \
		case 11: return CALL(5); case 13: return CALL(6); case 15: return CALL(7); \
		default: return og_set_error(OG_E_WINDOW, "window %d is not one of the warp kernel's (3 .. 15, odd)", window); \
	}

int og_launch_sauvola_warp_page(const uint8_t *src, int h, int w, int64_t src_pitch, int window, double k, double r,
	uint8_t *dst, int64_t dst_pitch, cudaStream_t st) {
#define OG_WS_PAGE(R) ws_launch_page<R>(src, h, w, src_pitch, dst, dst_pitch, k, r, st)
	OG_WS_SWITCH(OG_WS_PAGE)
#undef OG_WS_PAGE
}

int og_launch_sauvola_warp_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int window, uint8_t *out, int64_t capacity, cudaStream_t st) {
#define OG_WS_STRIPS(R) ws_launch_strips<R>(gray, steps, pix_offsets, chunk_offsets, n, out, capacity, st)
	OG_WS_SWITCH(OG_WS_STRIPS)
#undef OG_WS_STRIPS
}
