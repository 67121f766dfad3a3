// Strip hand-off to the recogniser: the OCR stage keeps the strips that
// are at least 6 x 6 pixels and feeds them to its model in padded batches
// (origami/batch/detect/ocr.py:238-268: `if im.width >= min_width and im.height >= min_height`,
// then `self._predictor(images[i:i + chunk_size])`, whose input pipeline pads every line image of a
// batch to the widest one).  Here that assembly happens on the device, so the strips never travel
// to the host: packed strips -> one (n_kept, H, w_max) uint8 batch, right/bottom padded, plus the
// width of every kept strip and the index of the line it came from.
#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

// One CTA: keep-flags, exclusive scan, widths.  n is a page's worth of lines (a few thousand).
__global__ void __launch_bounds__(1024) og_pack_plan_kernel(const int32_t *__restrict__ steps, int n, int min_w, int min_h,
	int H, int w_max, int32_t *__restrict__ slot_out, int32_t *__restrict__ widths_out, int32_t *__restrict__ kept_index_out,
	int32_t *__restrict__ n_kept_out) {
	__shared__ int warp_sums[33];
	__shared__ int s_carry;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	if (threadIdx.x == 0) s_carry = 0;
	__syncthreads();
	for (int base = 0; base < n; base += blockDim.x) {
		const int i = base + threadIdx.x;
		int keep = 0, xs = 0;
		if (i < n) {
			xs = steps[2 * i];
			const int ys = steps[2 * i + 1];
			// the reference's filter, plus what the fixed batch shape cannot hold
			keep = xs >= min_w && ys >= min_h && ys <= H && xs <= w_max;
		}
		int inc = keep;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const int t = __shfl_up_sync(0xffffffffu, inc, o);
			if (lane >= o) inc += t;
		}
		if (lane == 31) warp_sums[wid] = inc;
		__syncthreads();
		if (wid == 0) {
			int w = lane < (blockDim.x >> 5) ? warp_sums[lane] : 0, winc = w;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const int t = __shfl_up_sync(0xffffffffu, winc, o);
				if (lane >= o) winc += t;
			}
			warp_sums[lane] = winc - w;
			if (lane == 31) warp_sums[32] = winc;
		}
		__syncthreads();
		const int carry = s_carry;
		const int slot = carry + warp_sums[wid] + inc - keep;
		if (i < n) {
			slot_out[i] = keep ? slot : -1;
			if (keep) {
				widths_out[slot] = xs;
				kept_index_out[slot] = i;
			}
		}
		__syncthreads();
		if (threadIdx.x == 0) s_carry = carry + warp_sums[32];
		__syncthreads();
	}
	if (threadIdx.x == 0) *n_kept_out = s_carry;
}

// One CTA per line: rows of the strip are copied into the batch slot and padded to w_max; rows
// below the strip (ysteps < H) are padding too.
__global__ void __launch_bounds__(256) og_pack_copy_kernel(const uint8_t *__restrict__ strips, const int32_t *__restrict__ steps,
	const int64_t *__restrict__ pix_offsets, int n, int64_t capacity, const int32_t *__restrict__ slot_of, int H, int w_max, int pad_value,
	uint8_t *__restrict__ batch) {
	const unsigned int pad4 = (unsigned int)(pad_value & 255) * 0x01010101u;
	for (int li = blockIdx.x; li < n; li += gridDim.x) {
		const int slot = slot_of[li];
		if (slot < 0) continue;
		const int xs = steps[2 * li], ys = steps[2 * li + 1];
		const int64_t off = pix_offsets[li];
		const bool present = off + (int64_t)xs * ys <= capacity;  // a strip the extraction skipped: all padding
		const uint8_t *g = strips + off;
		uint8_t *o = batch + (int64_t)slot * H * w_max;
		const bool al = (w_max & 3) == 0 && (reinterpret_cast<uintptr_t>(o) & 3) == 0;
		if (al) {
			// 4 output bytes per thread and step; the source row is read byte-wise (packed strips have no row alignment)
			const int wq = w_max >> 2;
			for (int e = threadIdx.x; e < H * wq; e += blockDim.x) {
				const int r = e / wq, c = (e - r * wq) << 2;
				unsigned int v = pad4;
				if (present && r < ys && c < xs) {
					const uint8_t *s = g + (int64_t)r * xs + c;
					const int m = min(4, xs - c);
					unsigned int b = 0;
					for (int q = 0; q < 4; ++q) b |= (unsigned int)(q < m ? s[q] : (uint8_t)pad_value) << (8 * q);
					v = b;
				}
				reinterpret_cast<unsigned int *>(o)[e] = v;
			}
		} else {
			for (int e = threadIdx.x; e < H * w_max; e += blockDim.x) {
				const int r = e / w_max, c = e - r * w_max;
				o[e] = (present && r < ys && c < xs) ? g[(int64_t)r * xs + c] : (uint8_t)pad_value;
			}
		}
	}
}

// Binarized pixels as one bit each (numpy.packbits order: the first pixel is the top bit).  A thread
// packs 32 pixels: the low bit of every byte's OR-fold, gathered into nibbles by a multiplication
// whose partial products do not overlap (bytes b0..b3 at bits 0, 8, 16, 24 times
// 2^31 + 2^22 + 2^13 + 2^4 land on bits 31, 30, 29, 28).
__device__ __forceinline__ unsigned og_nibble(unsigned w) {
	w |= w >> 4;
	w |= w >> 2;
	w |= w >> 1;
	return ((w & 0x01010101u) * 0x80402010u) >> 28;
}

__global__ void __launch_bounds__(256) og_pack_bits_kernel(const uint8_t *__restrict__ src, int64_t n, uint8_t *__restrict__ dst) {
	const int64_t nout = (n + 7) >> 3;  // bytes
	const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 3) == 0);
	const int64_t ngroups = (nout + 3) >> 2;  // 4 output bytes = 32 pixels
	for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += (int64_t)gridDim.x * blockDim.x) {
		const int64_t p0 = g * 32;
		if (vec && p0 + 32 <= n) {
			const uint4 a = __ldg(reinterpret_cast<const uint4 *>(src + p0)), b = __ldg(reinterpret_cast<const uint4 *>(src + p0 + 16));
			const unsigned o0 = og_nibble(a.x) << 4 | og_nibble(a.y), o1 = og_nibble(a.z) << 4 | og_nibble(a.w);
			const unsigned o2 = og_nibble(b.x) << 4 | og_nibble(b.y), o3 = og_nibble(b.z) << 4 | og_nibble(b.w);
			*reinterpret_cast<unsigned *>(dst + g * 4) = o0 | o1 << 8 | o2 << 16 | o3 << 24;
		} else {
			for (int k = 0; k < 4 && g * 4 + k < nout; ++k) {
				unsigned o = 0;
				for (int j = 0; j < 8; ++j) {
					const int64_t q = p0 + 8 * k + j;
					o |= (q < n && src[q] != 0 ? 1u : 0u) << (7 - j);
				}
				dst[g * 4 + k] = (uint8_t)o;
			}
		}
	}
}

}  // namespace

extern "C" int og_pack_strips(const uint8_t *strips, const int32_t *steps, const int64_t *pix_offsets, int n, int64_t capacity,
	int min_w, int min_h, int pad_value, int H, int w_max,
	uint8_t *batch_out, int32_t *slot_scratch, int32_t *widths_out, int32_t *kept_index_out, int32_t *n_kept_out, void *stream) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	OG_REQUIRE(H > 0 && w_max > 0 && min_w >= 0 && min_h >= 0, OG_E_ARG, "batch shape (%d, %d) / minimum size invalid", H, w_max);
	OG_REQUIRE(n_kept_out != nullptr, OG_E_ARG, "n_kept_out is NULL");
	cudaStream_t st = (cudaStream_t)stream;
	if (n == 0) {
		cudaError_t e = cudaMemsetAsync(n_kept_out, 0, sizeof(int32_t), st);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
		return OG_OK;
	}
	OG_REQUIRE(strips != nullptr && steps != nullptr && pix_offsets != nullptr && batch_out != nullptr && slot_scratch != nullptr &&
		widths_out != nullptr && kept_index_out != nullptr, OG_E_ARG, "NULL pointer");
	og_pack_plan_kernel<<<1, 1024, 0, st>>>(steps, n, min_w, min_h, H, w_max, slot_scratch, widths_out, kept_index_out, n_kept_out);
	int rc = og_check_launch("og_pack_plan_kernel");
	if (rc) return rc;
	og_pack_copy_kernel<<<min(n, og_num_sms() * 8), 256, 0, st>>>(strips, steps, pix_offsets, n, capacity, slot_scratch, H, w_max, pad_value, batch_out);
	return og_check_launch("og_pack_copy_kernel");
}

extern "C" int og_pack_bits(const uint8_t *src, int64_t n, uint8_t *dst, void *stream) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	if (n == 0) return OG_OK;
	OG_REQUIRE(src != nullptr && dst != nullptr, OG_E_ARG, "NULL pointer");
	const int64_t groups = (n + 31) / 32;
	const int64_t want = (groups + 255) / 256;
	const int blocks = (int)(want < (int64_t)og_num_sms() * 16 ? want : (int64_t)og_num_sms() * 16);
	og_pack_bits_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, n, dst);
	return og_check_launch("og_pack_bits_kernel");
}
