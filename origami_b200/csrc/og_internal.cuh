// Declarations shared between the translation units of liborigami_b200 (not part of the C ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/origami_b200.h"

// Ordinal of the current CUDA device (-1 on error) and its SM count (cached per device).
int og_device_ordinal();
int og_num_sms();

// Per-device "already configured" flags for cudaFuncSetAttribute (slot = a small integer a kernel
// picks for itself).  Returns true the first time (device, slot) is seen.
bool og_first_use(int slot);

// og_lines.cu: the engines that write warped grids, bounding boxes and label counts, and the
// strip engines of round 1 (float64 map-gather, float32-screened map-gather, staged float64).
int og_extract_lines_engines(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *strips_out, int64_t capacity, int32_t *bbox_out, cudaStream_t st);

// og_strips.cu: line-owning CTAs, binarizer fused.
int og_launch_strips_fused(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n, const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *out, uint8_t *gray, int64_t capacity, int otsu, int32_t *thresholds_out, cudaStream_t st);

// og_binarize.cu
int og_binarize_strips_impl(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int mode, int window, uint8_t *out, int64_t capacity,
	int32_t *thresholds_out, cudaStream_t st);

// og_sauvola_warp.cu: warp-autonomous Sauvola for odd windows 3 .. 15 (pages and packed strips)
int og_launch_sauvola_warp_page(const uint8_t *src, int h, int w, int64_t src_pitch, int window, double k, double r,
	uint8_t *dst, int64_t dst_pitch, cudaStream_t st);
int og_launch_sauvola_warp_strips(const uint8_t *gray, const int32_t *steps, const int64_t *pix_offsets,
	const int32_t *chunk_offsets, int n, int window, uint8_t *out, int64_t capacity, cudaStream_t st);

#ifdef __CUDACC__
// Four bytes at once: 0xFF where the byte of x exceeds t (0 .. 255), else 0.  x > t is the carry out of
// bit 7 of x + (255 - t): the low seven bits are added per byte without crossing into the neighbour,
// the carry out is the majority of the two top bits and that carry in, and prmt's sign-replicating
// selectors spread it over the byte.  Four instructions per word (a compare-select-insert per byte is ~6
// per pixel).  og_gt4_consts gives the two per-threshold constants.
struct og_gt4_t {
	unsigned c7, c8;
};
__device__ __forceinline__ og_gt4_t og_gt4_consts(unsigned t) {
	const unsigned c = 255u - t;
	og_gt4_t k;
	k.c7 = (c & 0x7Fu) * 0x01010101u;
	k.c8 = (c & 0x80u) * 0x01010101u;
	return k;
}
__device__ __forceinline__ unsigned og_gt4(unsigned x, og_gt4_t k) {
	const unsigned z = (x & 0x7F7F7F7Fu) + k.c7;
	const unsigned o = (x & k.c8) | ((x | k.c8) & z);
	unsigned d;
	asm("prmt.b32 %0, %1, %1, 0xBA98;" : "=r"(d) : "r"(o));
	return d;
}

// Otsu threshold from an exact 256-bin histogram (skimage 0.16.2 threshold_otsu on uint8: bins =
// the integers [min, max], float64 cumulative sums, first argmax of w1 * w2 * (m1 - m2)^2; a
// single-valued image is the ValueError -> 128 case of origami/core/binarize.py:20-23).  One
// warp; each lane owns 8 consecutive bins.
__device__ inline int og_otsu_from_hist_warp(const unsigned int *hist) {
	const int lane = threadIdx.x & 31;
	unsigned long long cnt = 0, sum = 0;
	int lo = 256, hi = -1;
	for (int k = 0; k < 8; ++k) {
		const int b = lane * 8 + k;
		const unsigned int h = hist[b];
		cnt += h;
		sum += (unsigned long long)h * b;
		if (h) {
			lo = min(lo, b);
			hi = max(hi, b);
		}
	}
	unsigned long long icnt = cnt, isum = sum;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned long long tc = __shfl_up_sync(0xffffffffu, icnt, o);
		unsigned long long ts = __shfl_up_sync(0xffffffffu, isum, o);
		if (lane >= o) {
			icnt += tc;
			isum += ts;
		}
	}
	const unsigned long long total_cnt = __shfl_sync(0xffffffffu, icnt, 31);
	const unsigned long long total_sum = __shfl_sync(0xffffffffu, isum, 31);
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
		hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
	}
	if (hi <= lo) return 128;  // empty or single colour
	// candidate thresholds i in [lo, hi - 1]: class 1 = bins <= i, class 2 = bins > i
	unsigned long long c1 = icnt - cnt, s1 = isum - sum;  // exclusive prefix
	double best = -1.0;
	int best_i = 0x7fffffff;
	for (int k = 0; k < 8; ++k) {
		const int b = lane * 8 + k;
		const unsigned int h = hist[b];
		c1 += h;
		s1 += (unsigned long long)h * b;
		if (b >= lo && b < hi) {
			const unsigned long long c2 = total_cnt - c1, s2 = total_sum - s1;
			const double w1 = (double)c1, w2 = (double)c2;
			const double m1 = __ddiv_rn((double)s1, w1), m2 = __ddiv_rn((double)s2, w2);
			const double d = __dsub_rn(m1, m2);
			const double v = __dmul_rn(__dmul_rn(w1, w2), __dmul_rn(d, d));
			if (v > best) {
				best = v;
				best_i = b;
			}
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const double ob = __shfl_xor_sync(0xffffffffu, best, o);
		const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
		if (ob > best || (ob == best && oi < best_i)) {
			best = ob;
			best_i = oi;
		}
	}
	return best_i;
}
#endif
