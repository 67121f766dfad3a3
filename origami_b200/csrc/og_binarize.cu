// Binarizers (A12): Otsu and Sauvola thresholds, on whole pages and on packed line strips.
//
// Reference: origami/core/binarize.py:11-43 (binarize_with_threshold: (pixels > t) * 255),
// called per line by Line.image (origami/core/block.py:140-141), on the page by
// Page.binarized (origami/core/page.py:101-104) and Regions.binarized
// (origami/batch/detect/layout.py:157-169).  The thresholds are scikit-image 0.16.2's
// threshold_otsu / threshold_sauvola; both are restated with exact integer window sums /
// histograms followed by the same float64 operations in the same order.
#include <cstdlib>
#include <mutex>

#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

constexpr int BTHREADS = 256;
constexpr int ST_W = 128;  // Sauvola output tile on pages: 32 rows x 128 columns
constexpr int ST_H = 32;
constexpr int MAX_WINDOW = 127;

// np.pad(mode="reflect") index (no edge repeat), any overhang
__device__ __forceinline__ int reflect_index(int i, int n) {
	if (n == 1) return 0;
	const int period = 2 * (n - 1);
	i %= period;
	if (i < 0) i += period;
	return i >= n ? period - i : i;
}

// same map, but the common cases (inside, or one reflection) avoid the integer division
__device__ __forceinline__ int reflect_index_fast(int i, int n) {
	if ((unsigned)i < (unsigned)n) return i;
	const int j = i < 0 ? -i : 2 * (n - 1) - i;
	if ((unsigned)j < (unsigned)n) return j;
	return reflect_index(i, n);
}

// ---------------------------------------------------------------------------------------
// strips, Otsu: one CTA per line -- histogram in shared memory, threshold, apply
// ---------------------------------------------------------------------------------------
// Histogram of text strips: most pixels are paper, so a warp's 32 lanes mostly hit the same few
// bins and plain shared-memory atomics serialise.  The counters are replicated HREP (4) times
// ([bin][replica], replica = lane % HREP), so equal values from different lanes land on
// different addresses and banks; the replicas are summed before the threshold is computed.
#ifndef OG_HREP
#define OG_HREP 4
#endif
constexpr int HREP = OG_HREP;

#ifndef OG_OTSU_STRIP_THREADS
#define OG_OTSU_STRIP_THREADS 128
#endif
#ifndef OG_OTSU_STRIP_CTAS
#define OG_OTSU_STRIP_CTAS 16
#endif
// a small CTA per line, all lines of a page resident at once: the kernel is a chain of short
// dependent phases (histogram, threshold, apply) and not bound by the shared-memory atomics
// (replacing them by plain reads does not change its time: 41 us for 2000 strips)
constexpr int OTHREADS = OG_OTSU_STRIP_THREADS;
#ifndef OG_OTSU_STRIP_UNROLL
#define OG_OTSU_STRIP_UNROLL 2
#endif
constexpr int OU = OG_OTSU_STRIP_UNROLL;  // 16-byte vectors in flight per thread

__global__ void __launch_bounds__(OTHREADS) og_otsu_strips_kernel(
	const uint8_t *__restrict__ gray, const int32_t *__restrict__ steps, const int64_t *__restrict__ pix_offsets,
	int n, uint8_t *__restrict__ out, int64_t capacity, int32_t *__restrict__ thresholds) {
	__shared__ unsigned int rep[256 * HREP];
	__shared__ unsigned int hist[256];
	__shared__ int s_t;
	const int r = threadIdx.x & (HREP - 1);
	for (int li = blockIdx.x; li < n; li += gridDim.x) {
		const int64_t off = pix_offsets[li];
		const int64_t npix = (int64_t)steps[2 * li] * steps[2 * li + 1];
		if (off + npix > capacity) continue;  // as og_extract_lines: a strip that does not fit is skipped whole
		const uint8_t *g = gray + off;
		uint8_t *o = out + off;
		for (int k = threadIdx.x; k < 256 * HREP; k += OTHREADS) rep[k] = 0;
		__syncthreads();
		// 16-byte vector body between the unaligned head and tail
		const int64_t head = min(npix, (int64_t)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15));
		const int64_t nvec = (npix - head) / 16;
		for (int64_t i = threadIdx.x; i < head; i += OTHREADS) atomicAdd(&rep[g[i] * HREP + r], 1u);
		const uint4 *gv = reinterpret_cast<const uint4 *>(g + head);
		// OU vectors per thread are requested before the first one is counted: the shared-memory
		// atomics keep the compiler from overlapping the loads of successive iterations itself
		for (int64_t i0 = threadIdx.x; i0 < nvec; i0 += OTHREADS * OU) {
			uint4 v[OU];
#pragma unroll
			for (int u = 0; u < OU; ++u) v[u] = i0 + u * OTHREADS < nvec ? __ldg(gv + i0 + u * OTHREADS) : make_uint4(0, 0, 0, 0);
#pragma unroll
			for (int u = 0; u < OU; ++u) {
				if (i0 + u * OTHREADS >= nvec) break;
				const unsigned int wds[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
				for (int q = 0; q < 4; ++q) {
#pragma unroll
					for (int b = 0; b < 4; ++b) atomicAdd(&rep[__byte_perm(wds[q], 0, 0x4440 + b) * HREP + r], 1u);
				}
			}
		}
		for (int64_t i = head + nvec * 16 + threadIdx.x; i < npix; i += OTHREADS) atomicAdd(&rep[g[i] * HREP + r], 1u);
		__syncthreads();
		for (int k = threadIdx.x; k < 256; k += OTHREADS) {
			unsigned int sum = 0;
#pragma unroll
			for (int q = 0; q < HREP; ++q) sum += rep[k * HREP + q];
			hist[k] = sum;
		}
		__syncthreads();
		if (threadIdx.x < 32) {
			const int t = og_otsu_from_hist_warp(hist);
			if (threadIdx.x == 0) {
				s_t = t;
				if (thresholds != nullptr) thresholds[li] = t;
			}
		}
		__syncthreads();
		const unsigned int t = (unsigned int)s_t;
		const og_gt4_t gt = og_gt4_consts(t);
		for (int64_t i = threadIdx.x; i < head; i += OTHREADS) o[i] = g[i] > t ? 255 : 0;
		const bool out_al = ((reinterpret_cast<uintptr_t>(o + head)) & 15) == 0;
		for (int64_t i0 = threadIdx.x; i0 < nvec; i0 += OTHREADS * OU) {
			uint4 v[OU];
#pragma unroll
			for (int u = 0; u < OU; ++u) v[u] = i0 + u * OTHREADS < nvec ? __ldg(gv + i0 + u * OTHREADS) : make_uint4(0, 0, 0, 0);
#pragma unroll
			for (int u = 0; u < OU; ++u) {
				const int64_t i = i0 + u * OTHREADS;
				if (i >= nvec) break;
				const unsigned int wds[4] = {og_gt4(v[u].x, gt), og_gt4(v[u].y, gt), og_gt4(v[u].z, gt), og_gt4(v[u].w, gt)};
				if (out_al) {
					reinterpret_cast<uint4 *>(o + head)[i] = make_uint4(wds[0], wds[1], wds[2], wds[3]);
				} else {
					uint8_t *ob = o + head + i * 16;
#pragma unroll
					for (int q = 0; q < 16; ++q) ob[q] = (uint8_t)(wds[q >> 2] >> (8 * (q & 3)));
				}
			}
		}
		for (int64_t i = head + nvec * 16 + threadIdx.x; i < npix; i += OTHREADS) o[i] = g[i] > t ? 255 : 0;
		__syncthreads();
	}
}

// ---------------------------------------------------------------------------------------
// page, Otsu: global histogram (scratch), then threshold + apply
// ---------------------------------------------------------------------------------------
// A CTA walks whole rows (grid-stride); a row is read as 16-byte vectors between an unaligned
// head and tail, so no per-pixel division is needed and loads are 128 bit.
__global__ void __launch_bounds__(BTHREADS) og_hist_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t pitch, unsigned int *__restrict__ hist_out) {
	__shared__ unsigned int rep[256 * HREP];
	for (int k = threadIdx.x; k < 256 * HREP; k += BTHREADS) rep[k] = 0;
	__syncthreads();
	const int r = threadIdx.x & (HREP - 1);
	for (int y = blockIdx.x; y < h; y += gridDim.x) {
		const uint8_t *g = src + (int64_t)y * pitch;
		const int head = min(w, (int)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15));
		const int nvec = (w - head) / 16;
		for (int i = threadIdx.x; i < head; i += BTHREADS) atomicAdd(&rep[g[i] * HREP + r], 1u);
		const uint4 *gv = reinterpret_cast<const uint4 *>(g + head);
		for (int i = threadIdx.x; i < nvec; i += BTHREADS) {
			const uint4 v = __ldg(gv + i);
			const unsigned int wds[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
			for (int q = 0; q < 4; ++q) {
#pragma unroll
				for (int b = 0; b < 4; ++b) atomicAdd(&rep[((wds[q] >> (8 * b)) & 255u) * HREP + r], 1u);
			}
		}
		for (int i = head + nvec * 16 + threadIdx.x; i < w; i += BTHREADS) atomicAdd(&rep[g[i] * HREP + r], 1u);
	}
	__syncthreads();
	for (int k = threadIdx.x; k < 256; k += BTHREADS) {
		unsigned int sum = 0;
#pragma unroll
		for (int q = 0; q < HREP; ++q) sum += rep[k * HREP + q];
		if (sum) atomicAdd(hist_out + k, sum);
	}
}

__global__ void __launch_bounds__(BTHREADS) og_otsu_apply_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t src_pitch, uint8_t *__restrict__ dst, int64_t dst_pitch,
	const unsigned int *__restrict__ hist_in, int32_t *__restrict__ threshold_out) {
	__shared__ unsigned int hist[256];
	__shared__ int s_t;
	for (int k = threadIdx.x; k < 256; k += BTHREADS) hist[k] = hist_in[k];
	__syncthreads();
	if (threadIdx.x < 32) {
		const int t = og_otsu_from_hist_warp(hist);
		if (threadIdx.x == 0) {
			s_t = t;
			if (blockIdx.x == 0 && threshold_out != nullptr) *threshold_out = t;
		}
	}
	__syncthreads();
	const unsigned int t = (unsigned int)s_t;
	const og_gt4_t gt = og_gt4_consts(t);
	for (int y = blockIdx.x; y < h; y += gridDim.x) {
		const uint8_t *g = src + (int64_t)y * src_pitch;
		uint8_t *o = dst + (int64_t)y * dst_pitch;
		const int head = min(w, (int)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15));
		const int nvec = (w - head) / 16;
		const bool out_al = ((reinterpret_cast<uintptr_t>(o + head)) & 15) == 0;
		for (int i = threadIdx.x; i < head; i += BTHREADS) o[i] = g[i] > t ? 255 : 0;
		const uint4 *gv = reinterpret_cast<const uint4 *>(g + head);
		for (int i = threadIdx.x; i < nvec; i += BTHREADS) {
			const uint4 v = __ldg(gv + i);
			const unsigned int wds[4] = {og_gt4(v.x, gt), og_gt4(v.y, gt), og_gt4(v.z, gt), og_gt4(v.w, gt)};
			if (out_al) {
				__stcs(reinterpret_cast<uint4 *>(o + head) + i, make_uint4(wds[0], wds[1], wds[2], wds[3]));
			} else {
				uint8_t *ob = o + head + i * 16;
#pragma unroll
				for (int q = 0; q < 16; ++q) ob[q] = (uint8_t)(wds[q >> 2] >> (8 * (q & 3)));
			}
		}
		for (int i = head + nvec * 16 + threadIdx.x; i < w; i += BTHREADS) o[i] = g[i] > t ? 255 : 0;
	}
}

// ---------------------------------------------------------------------------------------
// Sauvola on one ST_H x ST_W output tile of an image (skimage 0.16.2 threshold_sauvola:
// reflect-padded centred window, m = S1/w^2, s = sqrt(max(S2/w^2 - m*m, 0)),
// t = m * (1 + k * ((s / r) - 1)) in float64).  Window sums are exact integers.
// ---------------------------------------------------------------------------------------
template <int TW>
__device__ void sauvola_tile(const uint8_t *__restrict__ img, int h, int w, int64_t pitch,
	uint8_t *__restrict__ out, int64_t out_pitch, int ty0, int tx0, int win, double kk, double rr,
	unsigned char *smem) {
	const int rad = win / 2;
	const int PW = TW + 2 * rad, PH = ST_H + 2 * rad;
	const int PV = PW | 1;  // odd row stride: lanes that own different rows hit different banks
	// The halo tile is stored with a left margin of XO >= rad columns, XO a multiple of 16, so that
	// a tile row starts at a 16-byte aligned source address (tx0 is a multiple of TW) and is filled
	// with 16-byte loads; padded column px lives at tile[py * TP + px].
	const int XO = (rad + 15) & ~15;
	const int TPITCH = TW + 2 * XO;
	uint8_t *tile0 = smem;                                         // PH x TPITCH halo tile
	uint8_t *tile = tile0 + (XO - rad);
	const int TP = TPITCH;
	unsigned int *v1 = reinterpret_cast<unsigned int *>(smem + ((PH * TPITCH + 15) & ~15));  // ST_H x PV vertical sums of p
	unsigned int *v2 = v1 + ST_H * PV;                                                   // ... of p^2
	uint8_t *otile = reinterpret_cast<uint8_t *>(v2 + ST_H * PV);  // ST_H x (TW + 16) results
	const int tid = threadIdx.x;
	const int wid = tid >> 5, lane = tid & 31;
	__syncthreads();
	{
		// halo tile: a warp per padded row, lanes along x (no per-element division); only tiles
		// that touch the image border pay for the reflect-101 index arithmetic
		const bool interior = ty0 - rad >= 0 && ty0 + ST_H + rad <= h && tx0 - rad >= 0 && tx0 + TW + rad <= w;
		const bool vec = ty0 - rad >= 0 && ty0 + ST_H + rad <= h && tx0 - XO >= 0 && tx0 + TW + XO <= w &&
			((reinterpret_cast<uintptr_t>(img) | (uintptr_t)pitch | (uintptr_t)tx0) & 15) == 0;
		if (vec) {
			// whole margins inside the image: 16 bytes per thread and load
			const int CH = TPITCH >> 4;
			const uint8_t *base = img + (int64_t)(ty0 - rad) * pitch + (tx0 - XO);
			for (int c = tid; c < PH * CH; c += BTHREADS) {
				const int py = c / CH, cx = (c - py * CH) << 4;
				*reinterpret_cast<uint4 *>(tile0 + py * TPITCH + cx) = __ldg(reinterpret_cast<const uint4 *>(base + (int64_t)py * pitch + cx));
			}
		} else {
			for (int py = wid; py < PH; py += BTHREADS / 32) {
				if (interior) {
					const uint8_t *rowp = img + (int64_t)(ty0 - rad + py) * pitch + (tx0 - rad);
					for (int px = lane; px < PW; px += 32) tile[py * TP + px] = __ldg(rowp + px);
				} else {
					const uint8_t *rowp = img + (int64_t)reflect_index_fast(ty0 - rad + py, h) * pitch;
					for (int px = lane; px < PW; px += 32) tile[py * TP + px] = __ldg(rowp + reflect_index_fast(tx0 - rad + px, w));
				}
			}
		}
	}
	__syncthreads();
	// vertical running sums: a thread owns a padded column and a band of output rows
	{
		const int groups = max(1, min(BTHREADS / PW, ST_H));
		const int g = tid / PW, px0 = tid - g * PW;
		const int band = (ST_H + groups - 1) / groups;
		const int pstep = (PW <= BTHREADS) ? PW * groups : BTHREADS;  // PW > BTHREADS: several passes
		if (g < groups) {
			const int ybeg = g * band, yend = min(ST_H, ybeg + band);
			for (int px = px0; px < PW; px += pstep) {
				unsigned int s1 = 0, s2 = 0;
				for (int py = ybeg; py < ybeg + win; ++py) {
					const unsigned int v = tile[py * TP + px];
					s1 += v;
					s2 += v * v;
				}
				v1[ybeg * PV + px] = s1;
				v2[ybeg * PV + px] = s2;
				for (int y = ybeg + 1; y < yend; ++y) {
					const unsigned int a = tile[(y - 1) * TP + px], b = tile[(y + win - 1) * TP + px];
					s1 += b - a;
					s2 += b * b - a * a;
					v1[y * PV + px] = s1;
					v2[y * PV + px] = s2;
				}
			}
		}
	}
	__syncthreads();
	// horizontal running sums + threshold: lane = row (conflict-free with the odd stride), warp =
	// a run of SEG consecutive columns, so the window slides SEG - 1 times per initial sum
	{
		constexpr int SEG = TW / (BTHREADS / 32);
		const int row = lane, xb = wid * SEG;
		const unsigned int *r1 = v1 + row * PV, *r2 = v2 + row * PV;
		unsigned int s1 = 0, s2 = 0;
		for (int j = 0; j < win; ++j) {
			s1 += r1[xb + j];
			s2 += r2[xb + j];
		}
		const double n = (double)(win * win);
		const unsigned int nwin = (unsigned int)(win * win);
		const bool small_win = nwin <= 256u;
		const float c1 = (float)((1.0 - kk) / n), c2 = (float)(kk / (n * n * rr));
		// constant trip count and pointers hoisted: every access below is base + immediate
		const unsigned int *in1 = r1 + xb + win - 1, *in2 = r2 + xb + win - 1;  // column entering the window
		const unsigned int *out1 = r1 + xb - 1, *out2 = r2 + xb - 1;            // column leaving it
		const uint8_t *centre = tile + (row + rad) * TP + xb + rad;
		uint8_t *res = otile + row * (TW + 16) + xb;
#pragma unroll
		for (int j = 0; j < SEG; ++j) {
			if (j > 0) {
				s1 += in1[j] - out1[j];
				s2 += in2[j] - out2[j];
			}
			const int pv = centre[j];
			// Screening in fp32: t = (A/n)(1 - k + k sqrt(nB - A^2)/(n r)) with the variance
			// numerator D = nB - A^2 formed exactly in 64-bit integers (no cancellation).  Its
			// error is < 2e-4 gray levels, so a pixel further than 1e-3 from the estimate is
			// decided by it; the few others take the float64 sequence of the reference.
			// (up to 15 x 15 windows n * S2 <= n^2 * 255^2 < 2^32: 32-bit arithmetic gives the same D)
			const float fD = small_win ? (float)(nwin * s2 - s1 * s1)
				: (float)((unsigned long long)nwin * s2 - (unsigned long long)s1 * s1);
			const float ta = __fmul_rn((float)s1, __fadd_rn(c1, __fmul_rn(c2, sqrtf(fD))));
			const float diff = __fsub_rn((float)pv, ta);
			bool white;
			if (fabsf(diff) > 1.0e-3f) {
				white = diff > 0.f;
			} else {
				const double m = __ddiv_rn((double)s1, n);
				const double g2 = __ddiv_rn((double)s2, n);
				double var = __dsub_rn(g2, __dmul_rn(m, m));
				var = var > 0.0 ? var : 0.0;
				const double sd = __dsqrt_rn(var);
				const double t = __dmul_rn(m, __dadd_rn(1.0, __dmul_rn(kk, __dsub_rn(__ddiv_rn(sd, rr), 1.0))));
				white = (double)pv > t;
			}
			res[j] = white ? 255 : 0;
		}
	}
	__syncthreads();
	// coalesced write-out: 16-byte chunks when the destination allows it, else one byte per lane
	{
		const int rows = min(ST_H, h - ty0), cols = min(TW, w - tx0);
		const bool al = ((reinterpret_cast<uintptr_t>(out) | (uintptr_t)out_pitch | (uintptr_t)tx0) & 15) == 0;
		if (al) {
			constexpr int CPR = TW / 16;
			for (int c = tid; c < ST_H * CPR; c += BTHREADS) {
				const int row = c / CPR, cx = (c % CPR) * 16;
				if (row >= rows || cx >= cols) continue;
				uint8_t *o = out + (int64_t)(ty0 + row) * out_pitch + tx0 + cx;
				const uint8_t *t = otile + row * (TW + 16) + cx;
				if (cx + 16 <= cols) {
					*reinterpret_cast<uint4 *>(o) = *reinterpret_cast<const uint4 *>(t);
				} else {
					for (int q = 0; cx + q < cols; ++q) o[q] = t[q];
				}
			}
		} else {
			for (int c = tid; c < ST_H * TW; c += BTHREADS) {
				const int row = c / TW, cx = c % TW;
				if (row < rows && cx < cols) out[(int64_t)(ty0 + row) * out_pitch + tx0 + cx] = otile[row * (TW + 16) + cx];
			}
		}
	}
}

template <int TW>
__host__ __device__ inline size_t sauvola_smem_bytes(int win) {
	const int rad = win / 2;
	const int PW = TW + 2 * rad, PH = ST_H + 2 * rad;
	const int TPITCH = TW + 2 * ((rad + 15) & ~15);
	return (size_t)((PH * TPITCH + 15) & ~15) + (size_t)2 * ST_H * (PW | 1) * sizeof(unsigned int) + (size_t)ST_H * (TW + 16);
}

__global__ void __launch_bounds__(BTHREADS) og_sauvola_page_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t src_pitch, uint8_t *__restrict__ dst, int64_t dst_pitch,
	int win, double kk, double rr) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int tiles_x = (w + ST_W - 1) / ST_W, tiles_y = (h + ST_H - 1) / ST_H;
	const int64_t total = (int64_t)tiles_x * tiles_y;
	for (int64_t t = blockIdx.x; t < total; t += gridDim.x) {
		const int ty = (int)(t / tiles_x), tx = (int)(t - (int64_t)ty * tiles_x);
		sauvola_tile<ST_W>(src, h, w, src_pitch, dst, dst_pitch, ty * ST_H, tx * ST_W, win, kk, rr, smem_raw);
	}
}

// ---------------------------------------------------------------------------------------
// Column-streaming Sauvola (round 2).  A CTA owns a stripe of SS_W output columns and a run of
// rows and walks down the image: a thread per input column (stripe + window radius on both sides)
// keeps the sums of p and p^2 over the window's rows in registers and updates them with one row
// entering and one leaving, so every source byte is read twice (entering, leaving; the second
// read hits L1) instead of once per overlapping halo tile; the column sums of SS_R consecutive rows
// go to shared memory, then every thread slides the horizontal window over four neighbouring
// outputs of one of those rows (16-byte shared loads) and applies the threshold test of
// sauvola_tile (fp32 screen, the reference's float64 sequence for the few pixels within 1e-3 of
// the threshold).  Same exact integer window sums, same results.
// ---------------------------------------------------------------------------------------
constexpr int SS_W = 128;    // output columns per stripe
constexpr int SS_R = 8;      // rows per barrier interval
constexpr int SS_T = 256;    // threads: SS_R rows x (SS_W / 4) quads in the horizontal phase
static_assert(SS_R * (SS_W / 4) == SS_T, "one thread per (row, quad)");

__host__ __device__ inline int ss_pitch(int win) { return (SS_W + win - 1 + 3 + 4) & ~3; }  // u32 entries per row of column sums

template <bool QUAD_OK>
__device__ void sauvola_stream(const uint8_t *__restrict__ img, int h, int w, int64_t pitch,
	uint8_t *__restrict__ out, int64_t out_pitch, int tx0, int y0, int y1, int win, double kk, double rr,
	unsigned int *v1, unsigned int *v2) {
	const int rad = win / 2, ncol = SS_W + 2 * rad, NP = ss_pitch(win);
	const int tid = threadIdx.x;
	// ---- vertical state: thread t owns input column tx0 - rad + t (reflect-101 at the image border)
	// (ncol <= 128 + 126 = 254: one column per thread)
	// Interior stripes of 4-byte aligned images take the quad path: a thread owns four neighbouring
	// columns, reads them as two aligned words shifted into place, and writes its sums as 16-byte stores.
	const int xs0 = tx0 - rad;                        // image column of sum index 0
	const int nquad = (ncol + 3) >> 2;
	const bool quad = QUAD_OK && xs0 >= 0 && ((xs0 + 4 * nquad + 3) | 3) < (int)pitch && tx0 + SS_W + rad <= w &&
		((reinterpret_cast<uintptr_t>(img) | (uintptr_t)pitch) & 3) == 0;
	const bool has_col = quad ? tid < nquad : tid < ncol;
	const int qbyte = (xs0 + 4 * tid) & ~3, qshift = 8 * ((xs0 + 4 * tid) & 3);
	const uint8_t *colp = img + (quad ? qbyte : (has_col ? reflect_index_fast(xs0 + tid, w) : 0));
	unsigned int c1[4] = {0, 0, 0, 0}, c2[4] = {0, 0, 0, 0};
	auto quad_at = [&](int yy) {  // the four bytes of this thread's columns in image row yy
		const unsigned int *wp = reinterpret_cast<const unsigned int *>(colp + (int64_t)yy * pitch);
		return __funnelshift_r(wp[0], wp[1], qshift);
	};
	auto add_quad = [&](unsigned int v) {
#pragma unroll
		for (int j = 0; j < 4; ++j) {
			const unsigned int b = (v >> (8 * j)) & 255u;
			c1[j] += b;
			c2[j] += b * b;
		}
	};
	auto sub_quad = [&](unsigned int v) {
#pragma unroll
		for (int j = 0; j < 4; ++j) {
			const unsigned int b = (v >> (8 * j)) & 255u;
			c1[j] -= b;
			c2[j] -= b * b;
		}
	};
	if (has_col) {
		for (int yy = y0 - rad; yy < y0 + rad; ++yy) {  // rows y0 - rad .. y0 + rad - 1: the window of row y0 minus its last row
			const int ry = reflect_index_fast(yy, h);
			if (quad) add_quad(quad_at(ry));
			else {
				const unsigned int v = colp[(int64_t)ry * pitch];
				c1[0] += v;
				c2[0] += v * v;
			}
		}
	}
	// ---- horizontal role: row `hr` of the interval, outputs 4 * hq .. 4 * hq + 3 of the stripe
	const int hr = tid / (SS_W / 4), hq = tid % (SS_W / 4);
	const double n = (double)(win * win);
	const unsigned int nwin = (unsigned int)(win * win);
	const bool small_win = nwin <= 256u;
	const float fc1 = (float)((1.0 - kk) / n), fc2 = (float)(kk / (n * n * rr));
	const bool vec_io = ((reinterpret_cast<uintptr_t>(img) | (uintptr_t)pitch | reinterpret_cast<uintptr_t>(out) | (uintptr_t)out_pitch) & 3) == 0 && tx0 + SS_W <= w;

	for (int yb = y0; yb < y1; yb += SS_R) {
		__syncthreads();  // the previous interval's sums have been consumed
		if (has_col) {
			if (quad) {
#pragma unroll
				for (int i = 0; i < SS_R; ++i) {
					const int y = yb + i;
					add_quad(quad_at(reflect_index_fast(y + rad, h)));
					*reinterpret_cast<uint4 *>(v1 + i * NP + 4 * tid) = make_uint4(c1[0], c1[1], c1[2], c1[3]);
					*reinterpret_cast<uint4 *>(v2 + i * NP + 4 * tid) = make_uint4(c2[0], c2[1], c2[2], c2[3]);
					sub_quad(quad_at(reflect_index_fast(y - rad, h)));
				}
			} else {
#pragma unroll
				for (int i = 0; i < SS_R; ++i) {
					const int y = yb + i;
					const unsigned int vin = colp[(int64_t)reflect_index_fast(y + rad, h) * pitch];
					c1[0] += vin;
					c2[0] += vin * vin;
					v1[i * NP + tid] = c1[0];
					v2[i * NP + tid] = c2[0];
					const unsigned int vout = colp[(int64_t)reflect_index_fast(y - rad, h) * pitch];
					c1[0] -= vout;
					c2[0] -= vout * vout;
				}
			}
		}
		__syncthreads();
		const int y = yb + hr;
		if (y < y1 && y < h && tx0 + 4 * hq < w) {
			const unsigned int *r1 = v1 + hr * NP + 4 * hq, *r2 = v2 + hr * NP + 4 * hq;
			// window of output j covers column sums 4 hq + j .. 4 hq + j + 2 rad
			unsigned int s1 = 0, s2 = 0;
			int e = 0;
			for (; e + 4 <= win; e += 4) {
				const uint4 a = *reinterpret_cast<const uint4 *>(r1 + e), b = *reinterpret_cast<const uint4 *>(r2 + e);
				s1 += a.x + a.y + a.z + a.w;
				s2 += b.x + b.y + b.z + b.w;
			}
			for (; e < win; ++e) {
				s1 += r1[e];
				s2 += r2[e];
			}
			const uint8_t *cen = img + (int64_t)y * pitch + tx0 + 4 * hq;
			unsigned int pv4;
			const int nvalid = min(4, w - (tx0 + 4 * hq));
			if (vec_io) pv4 = *reinterpret_cast<const unsigned int *>(cen);
			else {
				pv4 = 0;
				for (int j = 0; j < nvalid; ++j) pv4 |= (unsigned int)cen[j] << (8 * j);
			}
			unsigned int res = 0;
#pragma unroll
			for (int j = 0; j < 4; ++j) {
				if (j > 0) {
					s1 += r1[j - 1 + win] - r1[j - 1];
					s2 += r2[j - 1 + win] - r2[j - 1];
				}
				const int pv = (pv4 >> (8 * j)) & 255;
				// screening in fp32 as in sauvola_tile: D = n S2 - S1^2 exact in integers
				const float fD = small_win ? (float)(nwin * s2 - s1 * s1)
					: (float)((unsigned long long)nwin * s2 - (unsigned long long)s1 * s1);
				// s1 < 2^23 and pv < 2^8: exact integer -> float through the mantissa of 2^23 (no conversion
				// instruction); the square root of the screen may be approximate (2 ulp): its error, 6e-5 gray
				// levels at most, stays far inside the 1e-3 margin that hands a pixel to the float64 sequence
				const float fs1 = __fsub_rn(__int_as_float(0x4B000000 | (int)s1), 8388608.0f);
				const float fpv = __fsub_rn(__int_as_float(0x4B000000 | pv), 8388608.0f);
				float rt;
				asm("sqrt.approx.f32 %0, %1;" : "=f"(rt) : "f"(fD));
				const float ta = __fmul_rn(fs1, __fadd_rn(fc1, __fmul_rn(fc2, rt)));
				const float diff = __fsub_rn(fpv, ta);
				bool white;
				if (fabsf(diff) > 1.0e-3f) {
					white = diff > 0.f;
				} else {
					const double m = __ddiv_rn((double)s1, n);
					const double g2 = __ddiv_rn((double)s2, n);
					double var = __dsub_rn(g2, __dmul_rn(m, m));
					var = var > 0.0 ? var : 0.0;
					const double sd = __dsqrt_rn(var);
					const double t = __dmul_rn(m, __dadd_rn(1.0, __dmul_rn(kk, __dsub_rn(__ddiv_rn(sd, rr), 1.0))));
					white = (double)pv > t;
				}
				res |= (white ? 255u : 0u) << (8 * j);
			}
			uint8_t *o = out + (int64_t)y * out_pitch + tx0 + 4 * hq;
			if (vec_io) *reinterpret_cast<unsigned int *>(o) = res;
			else for (int j = 0; j < nvalid; ++j) o[j] = (uint8_t)(res >> (8 * j));
		}
	}
}

__global__ void __launch_bounds__(SS_T) og_sauvola_stream_page_kernel(
	const uint8_t *__restrict__ src, int h, int w, int64_t src_pitch, uint8_t *__restrict__ dst, int64_t dst_pitch,
	int win, double kk, double rr, int seg_rows) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	unsigned int *v1 = reinterpret_cast<unsigned int *>(smem_raw);
	unsigned int *v2 = v1 + SS_R * ss_pitch(win);
	const int y0 = blockIdx.y * seg_rows;
	sauvola_stream<true>(src, h, w, src_pitch, dst, dst_pitch, blockIdx.x * SS_W, y0, min(h, y0 + seg_rows), win, kk, rr, v1, v2);
}

// strips: item = (line, stripe of SS_W columns); the strip is a (h, w) image of pitch w
__global__ void __launch_bounds__(SS_T) og_sauvola_stream_strips_kernel(
	const uint8_t *__restrict__ gray, const int32_t *__restrict__ steps, const int64_t *__restrict__ pix_offsets,
	int n, uint8_t *__restrict__ out, int64_t capacity, int win, double kk, double rr) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	unsigned int *v1 = reinterpret_cast<unsigned int *>(smem_raw);
	unsigned int *v2 = v1 + SS_R * ss_pitch(win);
	for (int li = blockIdx.x; li < n; li += gridDim.x) {
		const int w = steps[2 * li], h = steps[2 * li + 1];
		const int64_t off = pix_offsets[li];
		if (w <= 0 || h <= 0 || off + (int64_t)w * h > capacity) continue;  // a strip that does not fit is skipped whole
		for (int tx0 = 0; tx0 < w; tx0 += SS_W)
			sauvola_stream<false>(gray + off, h, w, w, out + off, w, tx0, 0, h, win, kk, rr, v1, v2);
	}
}

int check_window(int window) {
	OG_REQUIRE(window >= 3 && (window & 1) == 1, OG_E_WINDOW,
		"Window size w = %d must be odd and greater than 1.", window);
	OG_REQUIRE(window <= MAX_WINDOW, OG_E_WINDOW, "window %d larger than the supported maximum %d", window, MAX_WINDOW);
	return OG_OK;
}

}  // namespace

int og_binarize_strips_impl(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int mode, int window, uint8_t *out, int64_t capacity,
	int32_t *thresholds_out, cudaStream_t st) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	OG_REQUIRE(mode == OG_BIN_OTSU || mode == OG_BIN_SAUVOLA, OG_E_ARG, "mode %d unsupported", mode);
	if (n == 0) return OG_OK;
	OG_REQUIRE(gray != nullptr && out != nullptr && steps != nullptr && pix_offsets != nullptr && chunk_offsets != nullptr,
		OG_E_ARG, "NULL pointer");
	if (mode == OG_BIN_OTSU) {
		og_otsu_strips_kernel<<<min(n, og_num_sms() * OG_OTSU_STRIP_CTAS), OTHREADS, 0, st>>>(gray, steps, pix_offsets, n, out, capacity, thresholds_out);
		return og_check_launch("og_otsu_strips_kernel");
	}
	int rc = check_window(window);
	if (rc) return rc;
	OG_REQUIRE(gray != out, OG_E_ARG, "Sauvola cannot run in place");
	// windows <= 15: the warp kernel with ticketed items (289 us per page of strips against 363 us for the
	// column-streaming kernel below, which stays for larger windows and as OG_SAUVOLA_LEGACY)
	if (window <= 15 && getenv("OG_SAUVOLA_LEGACY") == nullptr)
		return og_launch_sauvola_warp_strips(gray, steps, pix_offsets, chunk_offsets, n, window, out, capacity, st);
	const size_t smem = (size_t)2 * SS_R * ss_pitch(window) * sizeof(unsigned int);
	og_sauvola_stream_strips_kernel<<<min(n, og_num_sms() * 6), SS_T, smem, st>>>(gray, steps, pix_offsets, n, out,
		capacity, window, 0.2, 127.5);
	return og_check_launch("og_sauvola_stream_strips_kernel");
}

extern "C" int og_binarize_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int mode, int window, uint8_t *out, int64_t capacity,
	int32_t *thresholds_out, void *stream) {
	return og_binarize_strips_impl(gray, steps, pix_offsets, chunk_offsets, n, mode, window, out, capacity,
		thresholds_out, (cudaStream_t)stream);
}

extern "C" int og_sauvola(const uint8_t *src, int h, int w, int64_t src_pitch, int window,
	double k, double r, uint8_t *dst, int64_t dst_pitch, void *stream) {
	OG_REQUIRE(src != nullptr && dst != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && src_pitch >= w && dst_pitch >= w, OG_E_ARG, "image %d x %d / pitch invalid", w, h);
	OG_REQUIRE(src != dst, OG_E_ARG, "Sauvola cannot run in place");
	int rc = check_window(window);
	if (rc) return rc;
	if (window <= 15 && getenv("OG_SAUVOLA_LEGACY") == nullptr)
		return og_launch_sauvola_warp_page(src, h, w, src_pitch, window, k, r, dst, dst_pitch, (cudaStream_t)stream);
	if (getenv("OG_SAUVOLA_STREAM") == nullptr) {
		// Pages: the tile kernel (0.160 ms for 4800x7200, window 15).  The column-streaming kernel below reads every
		// byte twice instead of once per overlapping halo tile, but its running sums serialise a CTA's rows
		// (0.21 - 0.25 ms measured): it is what the strips use (0.52 -> 0.37 ms) and stays selectable here.
		const size_t smem = sauvola_smem_bytes<ST_W>(window);
		if (og_first_use(56)) {  // once per device: the largest window's footprint
			cudaError_t e = cudaFuncSetAttribute(og_sauvola_page_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sauvola_smem_bytes<ST_W>(MAX_WINDOW));
			if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
		}
		const int64_t tiles = (int64_t)((w + ST_W - 1) / ST_W) * ((h + ST_H - 1) / ST_H);
		const int blocks = (int)(tiles < (int64_t)og_num_sms() * 8 ? tiles : (int64_t)og_num_sms() * 8);
		og_sauvola_page_kernel<<<blocks, BTHREADS, smem, (cudaStream_t)stream>>>(src, h, w, src_pitch, dst, dst_pitch, window, k, r);
		return og_check_launch("og_sauvola_page_kernel");
	}
	// stripes of SS_W columns x runs of rows, about six CTAs per SM; a run re-reads `window - 1` rows of warm-up
	const int stripes = (w + SS_W - 1) / SS_W;
	int segments = (og_num_sms() * 6 + stripes - 1) / stripes;
	segments = max(1, min(segments, (h + 4 * window - 1) / (4 * window)));  // runs of at least four window heights
	int seg_rows = (h + segments - 1) / segments;
	seg_rows = (seg_rows + SS_R - 1) / SS_R * SS_R;
	segments = (h + seg_rows - 1) / seg_rows;
	OG_REQUIRE(segments <= 65535, OG_E_RANGE, "image too tall");
	const size_t smem = (size_t)2 * SS_R * ss_pitch(window) * sizeof(unsigned int);
	og_sauvola_stream_page_kernel<<<dim3(stripes, segments), SS_T, smem, (cudaStream_t)stream>>>(src, h, w, src_pitch, dst, dst_pitch,
		window, k, r, seg_rows);
	return og_check_launch("og_sauvola_stream_page_kernel");
}

extern "C" int og_otsu(const uint8_t *src, int h, int w, int64_t src_pitch,
	uint8_t *dst, int64_t dst_pitch, uint32_t *hist_scratch, int32_t *threshold_out, void *stream) {
	OG_REQUIRE(src != nullptr && dst != nullptr && hist_scratch != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && src_pitch >= w && dst_pitch >= w, OG_E_ARG, "image %d x %d / pitch invalid", w, h);
	OG_REQUIRE((int64_t)h * w < ((int64_t)1 << 32), OG_E_RANGE, "image larger than 2^32 pixels");
	cudaStream_t st = (cudaStream_t)stream;
	cudaError_t e = cudaMemsetAsync(hist_scratch, 0, 256 * sizeof(uint32_t), st);
	if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
	const int blocks = og_num_sms() * 8;
	og_hist_kernel<<<blocks, BTHREADS, 0, st>>>(src, h, w, src_pitch, hist_scratch);
	int rc = og_check_launch("og_hist_kernel");
	if (rc) return rc;
	og_otsu_apply_kernel<<<blocks, BTHREADS, 0, st>>>(src, h, w, src_pitch, dst, dst_pitch, hist_scratch, threshold_out);
	return og_check_launch("og_otsu_apply_kernel");
}

// ---------------------------------------------------------------------------------------
// og_otsu_noalloc: page Otsu for callers that cannot obtain device memory.
//
// Who that is: an Origami stage plugin that shares its process with the segmentation
// networks.  TensorFlow reserves (nearly) the whole device for its own arena when it starts,
// so a later cudaMalloc from the plugin fails; the same holds for a worker that was handed a
// fixed set of buffers by a host application and has no allocator of its own, and for code
// on a latency-critical stream, where cudaMalloc / cudaFree would synchronise the device.
// Such a caller still owns its input and output images; what it lacks is the 1 KiB of
// histogram scratch that og_otsu expects from it.  Here that scratch lives in the library's
// own device image: a file-scope array that every GPU which loads the library holds until
// the process exits.
//
// Consequences of sharing one array:
//  - every call on a device uses the same 256 counters, and they keep the previous call's
//    counts, so each call clears them (cudaMemsetAsync on the caller's stream) before the
//    histogram kernel runs;
//  - two calls whose work overlaps on the GPU (different streams, or different host threads
//    without ordering between them) race on the counters and produce wrong thresholds:
//    serialise calls per device;
//  - the counters are 32 bit and the array has a fixed size, so the image size is capped
//    (OG_OTSU_NOALLOC_MAX_PIXELS); larger images are rejected with OG_E_RANGE.
//
// The array's name is only a host-side handle in host code, not a device address.  The
// first call on each device (with that device current) resolves the address once with
// cudaGetSymbolAddress and remembers it; make that first call outside any stream capture.
// The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
// ---------------------------------------------------------------------------------------
#define OG_OTSU_NOALLOC_MAX_PIXELS ((int64_t)1 << 30)
#define OG_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) uint32_t g_otsu_noalloc_hist[256];

namespace {

std::mutex g_noalloc_mutex;
uint32_t *g_noalloc_hist_ptr[OG_NOALLOC_MAX_DEVICES] = {nullptr};

int noalloc_hist_address(uint32_t **out) {
	int dev = -1;
	cudaError_t e = cudaGetDevice(&dev);
	if (e != cudaSuccess) return og_set_error((int)e, "cudaGetDevice: %s", cudaGetErrorString(e));
	OG_REQUIRE(dev >= 0 && dev < OG_NOALLOC_MAX_DEVICES, OG_E_RANGE, "device ordinal %d beyond %d", dev, OG_NOALLOC_MAX_DEVICES);
	std::lock_guard<std::mutex> lock(g_noalloc_mutex);
	if (g_noalloc_hist_ptr[dev] == nullptr) {
		void *ptr = nullptr;
		e = cudaGetSymbolAddress(&ptr, g_otsu_noalloc_hist);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaGetSymbolAddress: %s", cudaGetErrorString(e));
		g_noalloc_hist_ptr[dev] = static_cast<uint32_t *>(ptr);
	}
	*out = g_noalloc_hist_ptr[dev];
	return OG_OK;
}

}  // namespace

extern "C" int og_otsu_noalloc(const uint8_t *src, int h, int w, int64_t src_pitch,
	uint8_t *dst, int64_t dst_pitch, int32_t *threshold_out, void *stream) {
	OG_REQUIRE(src != nullptr && dst != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && src_pitch >= w && dst_pitch >= w, OG_E_ARG, "image %d x %d / pitch invalid", w, h);
	OG_REQUIRE((int64_t)h * w <= OG_OTSU_NOALLOC_MAX_PIXELS, OG_E_RANGE,
		"image of %lld pixels exceeds the %lld-pixel limit of the built-in scratch; use og_otsu with caller scratch",
		(long long)h * w, (long long)OG_OTSU_NOALLOC_MAX_PIXELS);
	uint32_t *hist = nullptr;
	int rc = noalloc_hist_address(&hist);
	if (rc) return rc;
	cudaStream_t st = (cudaStream_t)stream;
	cudaError_t e = cudaMemsetAsync(hist, 0, 256 * sizeof(uint32_t), st);  // the array keeps the last call's counts
	if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
	const int blocks = og_num_sms() * 8;
	og_hist_kernel<<<blocks, BTHREADS, 0, st>>>(src, h, w, src_pitch, hist);
	rc = og_check_launch("og_hist_kernel");
	if (rc) return rc;
	og_otsu_apply_kernel<<<blocks, BTHREADS, 0, st>>>(src, h, w, src_pitch, dst, dst_pitch, hist, threshold_out);
	return og_check_launch("og_otsu_apply_kernel");
}
