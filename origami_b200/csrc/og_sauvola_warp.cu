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
//   * a warp owns a stripe of 112 output columns and a run of rows; lane l holds the four pixels
//     4 l .. 4 l + 3 of the stripe's 128 input columns as one 32-bit word per row (one coalesced
//     128-byte row read per warp);
//   * horizontal: the neighbouring lanes' words come by shuffle (raw pixels, four shuffles), and the
//     window sums of p and p^2 of the lane's four outputs are dp4a dot products of those words
//     with byte masks (p: word . 0x01010101, p^2: (word & mask) . word) -- about seven
//     instructions per pixel for both sums;
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

constexpr int WS_OUT = 112;  // output columns per warp (lanes 2 .. 29)
constexpr int WS_WARPS = 4;  // warps per CTA (independent of each other)
#ifndef OG_WS_AHEAD
#define OG_WS_AHEAD 3
#endif
constexpr int WS_AHEAD = OG_WS_AHEAD;  // rows requested ahead of use

__device__ __forceinline__ int ws_reflect(int i, int n) {  // numpy.pad(mode="reflect"), any distance
	if ((unsigned)i < (unsigned)n) return i;
	if (n == 1) return 0;
	const int j = i < 0 ? -i : 2 * (n - 1) - i;
	if ((unsigned)j < (unsigned)n) return j;
	const int period = 2 * (n - 1);
	i %= period;
	if (i < 0) i += period;
	return i >= n ? period - i : i;
}

// bytes of the word of lane l + k (k = -2 .. 2) that lie in the window of the lane's output j
template <int RAD>
__device__ __forceinline__ constexpr unsigned ws_mask(int j, int k) {
	unsigned m = 0;
	for (int b = 0; b < 4; ++b) {
		const int d = 4 * k + b - j;
		if (d >= -RAD && d <= RAD) m |= 1u << (8 * b);
	}
	return m;
}

template <int RAD>
__device__ __forceinline__ constexpr unsigned ws_common(int k) {
	return ws_mask<RAD>(0, k) & ws_mask<RAD>(1, k) & ws_mask<RAD>(2, k) & ws_mask<RAD>(3, k);
}

// window sums of p (h1) and p^2 (h2) along the row for the lane's four outputs
template <int RAD>
__device__ __forceinline__ void ws_hsums(const unsigned (&wd)[5], unsigned (&h1)[4], unsigned (&h2)[4]) {
	unsigned b1 = 0, b2 = 0;
#pragma unroll
	for (int k = 0; k < 5; ++k) {
		const unsigned mc = ws_common<RAD>(k - 2);
		if (mc != 0) {
			b1 = __dp4a(wd[k], mc, b1);
			b2 = __dp4a(mc == 0x01010101u ? wd[k] : (wd[k] & (mc * 255u)), wd[k], b2);
		}
	}
#pragma unroll
	for (int j = 0; j < 4; ++j) {
		unsigned e1 = b1, e2 = b2;
#pragma unroll
		for (int k = 0; k < 5; ++k) {
			const unsigned m = ws_mask<RAD>(j, k - 2) & ~ws_common<RAD>(k - 2);
			if (m != 0) {
				e1 = __dp4a(wd[k], m, e1);
				e2 = __dp4a(wd[k] & (m * 255u), wd[k], e2);
			}
		}
		h1[j] = e1;
		h2[j] = e2;
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

// One stripe: output columns x0 .. x0 + 111, output rows y0 .. y1 - 1 of an (h, w) image
// (pitch and out_pitch below 2^31).
template <int RAD, bool ALIGNED>
__device__ __forceinline__ void ws_stripe(const uint8_t *__restrict__ img, int h, int w, unsigned pitch,
	uint8_t *__restrict__ out, unsigned out_pitch, int x0, int y0, int y1, const WsConst c,
	uint4 *__restrict__ ring_h, unsigned *__restrict__ ring_w) {
	constexpr int WIN = 2 * RAD + 1;
	constexpr unsigned NWIN = WIN * WIN;
	const int lane = threadIdx.x & 31;
	const int cx = x0 - 8 + 4 * lane;
	const bool inside = cx >= 0 && cx + 3 < w;
	const bool writes = lane >= 2 && lane < 30 && cx < w;
	const bool write_word = ALIGNED && writes && cx + 3 < w;
	const uint8_t *img_lane = img + (inside ? cx : 0);
	uint8_t *out_lane = out + cx;
	uint4 *rh = ring_h + lane;
	unsigned *rw = ring_w + lane;
#pragma unroll
	for (int i = 0; i < WIN; ++i) rh[i * 32] = make_uint4(0u, 0u, 0u, 0u);  // rows "before" the first leave as zeros
	unsigned v1[4] = {0, 0, 0, 0}, v2[4] = {0, 0, 0, 0};
	int slot = 0, cslot = WIN - RAD;  // ring slots of row r and of row r - RAD
	auto load_row = [&](int r) -> unsigned {
		const int ry = (unsigned)r < (unsigned)h ? r : ws_reflect(r, h);
		const uint8_t *row = img_lane + (uint64_t)(unsigned)ry * pitch;
		if (inside) {
			if (ALIGNED) return __ldg(reinterpret_cast<const unsigned *>(row));
			const uintptr_t a = reinterpret_cast<uintptr_t>(row);
			const unsigned *wp = reinterpret_cast<const unsigned *>(a & ~(uintptr_t)3);
			const unsigned sh = (unsigned)(a & 3) * 8;
			const unsigned lo = __ldg(wp);
			const unsigned hi = sh ? __ldg(wp + 1) : 0u;  // an aligned word that holds a byte of the image stays inside its page
			return __funnelshift_r(lo, hi, sh);
		}
		unsigned v = 0;
#pragma unroll
		for (int b = 0; b < 4; ++b) v |= (unsigned)__ldg(row + ws_reflect(cx + b, w)) << (8 * b);
		return v;
	};
	// rows are requested WS_AHEAD iterations before they are used (a warp has one 128-byte read per row
	// in flight otherwise, and nothing to do until it lands); a row past the end is a harmless reflection
	unsigned ahead[WS_AHEAD];
#pragma unroll
	for (int i = 0; i < WS_AHEAD; ++i) ahead[i] = load_row(y0 - RAD + i);
	for (int r = y0 - RAD; r < y1 + RAD; ++r) {
		unsigned wd[5];
		wd[2] = ahead[0];
#pragma unroll
		for (int i = 0; i + 1 < WS_AHEAD; ++i) ahead[i] = ahead[i + 1];
		ahead[WS_AHEAD - 1] = load_row(r + WS_AHEAD);
		wd[0] = __shfl_up_sync(0xffffffffu, wd[2], 2);
		wd[1] = __shfl_up_sync(0xffffffffu, wd[2], 1);
		wd[3] = __shfl_down_sync(0xffffffffu, wd[2], 1);
		wd[4] = __shfl_down_sync(0xffffffffu, wd[2], 2);
		unsigned h1[4], h2[4];
		ws_hsums<RAD>(wd, h1, h2);
		// the row entering the window replaces the row that left it WIN rows ago in the ring
		const uint4 old = rh[slot * 32];
		v1[0] += h1[0] - (old.x & 0xFFFu); v2[0] += h2[0] - (old.x >> 12);
		v1[1] += h1[1] - (old.y & 0xFFFu); v2[1] += h2[1] - (old.y >> 12);
		v1[2] += h1[2] - (old.z & 0xFFFu); v2[2] += h2[2] - (old.z >> 12);
		v1[3] += h1[3] - (old.w & 0xFFFu); v2[3] += h2[3] - (old.w >> 12);
		rh[slot * 32] = make_uint4(h2[0] * 4096u + h1[0], h2[1] * 4096u + h1[1], h2[2] * 4096u + h1[2], h2[3] * 4096u + h1[3]);
		rw[slot * 32] = wd[2];
		if (r >= y0 + RAD) {  // the window of output row r - RAD is complete
			const unsigned cw = rw[cslot * 32];
			float diff[4];
#pragma unroll
			for (int j = 0; j < 4; ++j) {
				// D = n S2 - S1^2 exactly (n^2 255^2 < 2^32 for windows <= 15); the float32 estimate of the
				// threshold is within 2e-4 gray levels of the float64 one (see sauvola_tile), and the
				// approximate square root adds 6e-5 at most: |pixel - estimate| > 1e-3 decides
				const unsigned D = NWIN * v2[j] - v1[j] * v1[j];
				float rt;
				asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(rt) : "f"((float)D));  // D is 0 or >= 1: nothing to flush
				const float fpv = __fsub_rn(__int_as_float((int)__byte_perm(cw, 0x4B000000u, 0x7650 + j)), 8388608.0f);
				diff[j] = __fmaf_rn(-(float)v1[j], __fmaf_rn(c.fc2, rt, c.fc1), fpv);
			}
			// bytes of sign bits: 0xFF where diff < 0 (or -0, which the slow path below settles)
			const unsigned n01 = ws_prmt(__float_as_uint(diff[0]), __float_as_uint(diff[1]), 0x00FBu);
			const unsigned n23 = ws_prmt(__float_as_uint(diff[2]), __float_as_uint(diff[3]), 0x00FBu);
			unsigned res = ~__byte_perm(n01, n23, 0x5410);
			const float closest = fminf(fminf(fabsf(diff[0]), fabsf(diff[1])), fminf(fabsf(diff[2]), fabsf(diff[3])));
			if (!(closest > 1.0e-3f)) {
#pragma unroll
				for (int j = 0; j < 4; ++j) {
					if (fabsf(diff[j]) > 1.0e-3f) continue;
					const double m = __ddiv_rn((double)v1[j], c.n);
					const double g2 = __ddiv_rn((double)v2[j], c.n);
					double var = __dsub_rn(g2, __dmul_rn(m, m));
					var = var > 0.0 ? var : 0.0;
					const double sd = __dsqrt_rn(var);
					const double t = __dmul_rn(m, __dadd_rn(1.0, __dmul_rn(c.kk, __dsub_rn(__ddiv_rn(sd, c.rr), 1.0))));
					const bool white = (double)((cw >> (8 * j)) & 255u) > t;
					res = (res & ~(255u << (8 * j))) | (white ? 255u << (8 * j) : 0u);
				}
			}
			uint8_t *o = out_lane + (uint64_t)(unsigned)(r - RAD) * out_pitch;
			if (write_word) {
				__stcs(reinterpret_cast<unsigned *>(o), res);
			} else if (writes) {
#pragma unroll
				for (int j = 0; j < 4; ++j)
					if (cx + j < w) o[j] = (uint8_t)(res >> (8 * j));
			}
		}
		slot = slot + 1 == WIN ? 0 : slot + 1;
		cslot = cslot + 1 == WIN ? 0 : cslot + 1;
	}
}

template <int RAD>
__host__ __device__ constexpr size_t ws_smem_bytes() {
	return (size_t)WS_WARPS * (2 * RAD + 1) * 32 * (sizeof(uint4) + sizeof(unsigned));
}

template <int RAD>
__device__ __forceinline__ void ws_rings(unsigned char *smem, uint4 *&ring_h, unsigned *&ring_w) {
	constexpr int WIN = 2 * RAD + 1;
	const int wid = threadIdx.x >> 5;
	ring_h = reinterpret_cast<uint4 *>(smem) + wid * WIN * 32;
	ring_w = reinterpret_cast<unsigned *>(smem + (size_t)WS_WARPS * WIN * 32 * sizeof(uint4)) + wid * WIN * 32;
}

// pages: item = (run of rows, stripe); neighbouring warps take neighbouring stripes of one run
template <int RAD, bool ALIGNED>
__global__ void __launch_bounds__(WS_WARPS * 32) og_sauvola_warp_page_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t src_pitch, uint8_t *__restrict__ dst, int64_t dst_pitch,
	WsConst c, int seg_rows, int stripes, int items) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	uint4 *ring_h;
	unsigned *ring_w;
	ws_rings<RAD>(smem_raw, ring_h, ring_w);
	const int item = blockIdx.x * WS_WARPS + (threadIdx.x >> 5);
	if (item >= items) return;
	const int seg = item / stripes, stripe = item - seg * stripes;
	const int y0 = seg * seg_rows;
	ws_stripe<RAD, ALIGNED>(src, h, w, (unsigned)src_pitch, dst, (unsigned)dst_pitch, stripe * WS_OUT, y0, min(h, y0 + seg_rows), c, ring_h, ring_w);
}

// strips: item = a 64-column chunk of the line plan; chunk q of a line works on stripe q when the
// line has that many stripes (a strip of w columns has ceil(w / 64) chunks >= ceil(w / 112) stripes)
template <int RAD>
__global__ void __launch_bounds__(WS_WARPS * 32) og_sauvola_warp_strips_kernel(
	const uint8_t *__restrict__ gray, const int32_t *__restrict__ steps, const int64_t *__restrict__ pix_offsets,
	const int32_t *__restrict__ chunk_offsets, int n, uint8_t *__restrict__ out, int64_t capacity, WsConst c) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	uint4 *ring_h;
	unsigned *ring_w;
	ws_rings<RAD>(smem_raw, ring_h, ring_w);
	const int total = __ldg(chunk_offsets + n);
	const int nwarps = gridDim.x * WS_WARPS;
	for (int item = blockIdx.x * WS_WARPS + (threadIdx.x >> 5); item < total; item += nwarps) {
		int lo = 0, hi = n;
		while (hi - lo > 1) {
			const int mid = (lo + hi) >> 1;
			if (__ldg(chunk_offsets + mid) <= item) lo = mid; else hi = mid;
		}
		const int q = item - __ldg(chunk_offsets + lo);
		const int w = __ldg(steps + 2 * lo), h = __ldg(steps + 2 * lo + 1);
		if (q * WS_OUT >= w || h <= 0) continue;
		const int64_t off = __ldg(pix_offsets + lo);
		if (off + (int64_t)w * h > capacity) continue;
		__syncwarp();
		ws_stripe<RAD, false>(gray + off, h, w, (unsigned)w, out + off, (unsigned)w, q * WS_OUT, 0, h, c, ring_h, ring_w);
		__syncwarp();
	}
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

template <int RAD>
int ws_launch_page(const uint8_t *src, int h, int w, int64_t src_pitch, uint8_t *dst, int64_t dst_pitch,
	double k, double r, cudaStream_t st) {
	const int stripes = (w + WS_OUT - 1) / WS_OUT;
	// About `per_sm` resident warps per SM in one wave (a run re-reads 2 RAD rows of warm-up, at half the
	// cost of an output row, so runs are kept long).
	static const int per_sm = getenv("OG_SAUVOLA_WARPS") ? atoi(getenv("OG_SAUVOLA_WARPS")) : 20;
	int segments = max(1, og_num_sms() * per_sm / stripes);
	segments = max(1, min(segments, (h + 4 * (2 * RAD + 1) - 1) / (4 * (2 * RAD + 1))));
	const int seg_rows = (h + segments - 1) / segments;
	segments = (h + seg_rows - 1) / seg_rows;
	const int64_t items64 = (int64_t)stripes * segments;
	OG_REQUIRE(items64 < ((int64_t)1 << 30) && src_pitch < ((int64_t)1 << 31) && dst_pitch < ((int64_t)1 << 31), OG_E_RANGE, "image too large");
	const int items = (int)items64;
	const int blocks = (items + WS_WARPS - 1) / WS_WARPS;
	const WsConst c = ws_const(2 * RAD + 1, k, r);
	const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)src_pitch | reinterpret_cast<uintptr_t>(dst) | (uintptr_t)dst_pitch) & 3) == 0;
	constexpr size_t smem = ws_smem_bytes<RAD>();
	if (aligned) {
		if (og_first_use(16 + 2 * RAD))
			cudaFuncSetAttribute(og_sauvola_warp_page_kernel<RAD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		og_sauvola_warp_page_kernel<RAD, true><<<blocks, WS_WARPS * 32, smem, st>>>(src, h, w, src_pitch, dst, dst_pitch, c, seg_rows, stripes, items);
	} else {
		if (og_first_use(17 + 2 * RAD))
			cudaFuncSetAttribute(og_sauvola_warp_page_kernel<RAD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		og_sauvola_warp_page_kernel<RAD, false><<<blocks, WS_WARPS * 32, smem, st>>>(src, h, w, src_pitch, dst, dst_pitch, c, seg_rows, stripes, items);
	}
	return og_check_launch("og_sauvola_warp_page_kernel");
}

template <int RAD>
int ws_launch_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, uint8_t *out, int64_t capacity, cudaStream_t st) {
	constexpr size_t smem = ws_smem_bytes<RAD>();
	if (og_first_use(32 + RAD))
		cudaFuncSetAttribute(og_sauvola_warp_strips_kernel<RAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	const WsConst c = ws_const(2 * RAD + 1, 0.2, 127.5);
	og_sauvola_warp_strips_kernel<RAD><<<og_num_sms() * 5, WS_WARPS * 32, smem, st>>>(gray, steps, pix_offsets, chunk_offsets, n, out, capacity, c);
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
