// Segmentation post-process (A18/A19) and the deskew affine warp (A11).
//
// Reference: label argmax origami/core/predict.py:281-285
// (np.argmax(np.sum(pr_masks, axis=0), -1).astype(uint8)), class masks
// origami/core/contours.py:488 (pixels == c.value), Line._extract_deskewed
// origami/core/block.py:155-172 (cv2.warpAffine, BORDER_CONSTANT).
// HBM-bound streaming kernels: 4*M*C bytes read and 1 + n_mask bytes written per pixel.
#include "og_common.cuh"

namespace {

constexpr int MAX_MASKS = 16;

struct MaskValues {
	uint8_t v[MAX_MASKS];
};

// C == 4: a thread owns 4 consecutive pixels -> float4 loads, uchar4 stores
__global__ void __launch_bounds__(256) og_argmax4_kernel(
	const float4 *__restrict__ probs, int M, int64_t npix,
	uint8_t *__restrict__ labels, uint8_t *__restrict__ masks, MaskValues mv, int n_mask) {
	const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const int64_t i0 = q * 4;
	if (i0 >= npix) return;
	const int cnt = (int)min((int64_t)4, npix - i0);
	uint32_t lab[4] = {0, 0, 0, 0};
#pragma unroll
	for (int k = 0; k < 4; ++k) {
		if (k < cnt) {
			float4 acc = __ldcs(probs + i0 + k);
			for (int m = 1; m < M; ++m) {
				const float4 v = __ldcs(probs + (int64_t)m * npix + i0 + k);
				acc.x = __fadd_rn(acc.x, v.x);
				acc.y = __fadd_rn(acc.y, v.y);
				acc.z = __fadd_rn(acc.z, v.z);
				acc.w = __fadd_rn(acc.w, v.w);
			}
			float best = acc.x;
			uint32_t bi = 0;
			if (acc.y > best) { best = acc.y; bi = 1; }
			if (acc.z > best) { best = acc.z; bi = 2; }
			if (acc.w > best) { best = acc.w; bi = 3; }
			lab[k] = bi;
		}
	}
	const bool vec = cnt == 4 && ((reinterpret_cast<uintptr_t>(labels) & 3) == 0);
	if (vec) {
		reinterpret_cast<uint32_t *>(labels)[q] = lab[0] | (lab[1] << 8) | (lab[2] << 16) | (lab[3] << 24);
	} else {
		for (int k = 0; k < cnt; ++k) labels[i0 + k] = (uint8_t)lab[k];
	}
	if (masks != nullptr) {
		const bool vecm = cnt == 4 && ((reinterpret_cast<uintptr_t>(masks) & 3) == 0) && ((npix & 3) == 0);
		for (int j = 0; j < n_mask; ++j) {
			const uint32_t val = mv.v[j];
			uint8_t *mrow = masks + (int64_t)j * npix;
			if (vecm) {
				reinterpret_cast<uint32_t *>(mrow)[q] = (uint32_t)(lab[0] == val) | ((uint32_t)(lab[1] == val) << 8) |
					((uint32_t)(lab[2] == val) << 16) | ((uint32_t)(lab[3] == val) << 24);
			} else {
				for (int k = 0; k < cnt; ++k) mrow[i0 + k] = lab[k] == val;
			}
		}
	}
}

// C == 4 with the model count known at compile time: every load of a thread is issued before the first
// sum (4 pixels x MT models = up to 20 independent 16-byte loads in flight per thread -- the runtime loop
// above waits for each model's load in turn and reaches a third of the HBM rate when its input is not in
// L2), and a lane's four pixels are 256 apart, so that every warp-wide load is 512 contiguous bytes.
template <int MT>
__global__ void __launch_bounds__(256) og_argmax4_unrolled_kernel(
	const float4 *__restrict__ probs, int64_t npix,
	uint8_t *__restrict__ labels, uint8_t *__restrict__ masks, MaskValues mv, int n_mask) {
	const int64_t base = (int64_t)blockIdx.x * 1024 + threadIdx.x;
	float4 v[MT][4];
#pragma unroll
	for (int m = 0; m < MT; ++m)
#pragma unroll
		for (int k = 0; k < 4; ++k) {
			const int64_t i = base + 256 * k;
			v[m][k] = i < npix ? __ldcs(probs + (int64_t)m * npix + i) : make_float4(0.f, 0.f, 0.f, 0.f);
		}
#pragma unroll
	for (int k = 0; k < 4; ++k) {
		const int64_t i = base + 256 * k;
		if (i >= npix) continue;
		float4 acc = v[0][k];
#pragma unroll
		for (int m = 1; m < MT; ++m) {  // np.sum(pr_masks, axis=0): in model order
			acc.x = __fadd_rn(acc.x, v[m][k].x);
			acc.y = __fadd_rn(acc.y, v[m][k].y);
			acc.z = __fadd_rn(acc.z, v[m][k].z);
			acc.w = __fadd_rn(acc.w, v[m][k].w);
		}
		float best = acc.x;
		uint32_t bi = 0;
		if (acc.y > best) { best = acc.y; bi = 1; }
		if (acc.z > best) { best = acc.z; bi = 2; }
		if (acc.w > best) { best = acc.w; bi = 3; }
		labels[i] = (uint8_t)bi;
		if (masks != nullptr)
			for (int j = 0; j < n_mask; ++j) masks[(int64_t)j * npix + i] = (bi == mv.v[j]);
	}
}

// any C: one pixel per thread
__global__ void __launch_bounds__(256) og_argmax_kernel(
	const float *__restrict__ probs, int M, int C, int64_t npix,
	uint8_t *__restrict__ labels, uint8_t *__restrict__ masks, MaskValues mv, int n_mask) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= npix) return;
	float best = 0.f;
	int bi = 0;
	for (int c = 0; c < C; ++c) {
		float acc = probs[i * C + c];
		for (int m = 1; m < M; ++m) acc = __fadd_rn(acc, probs[((int64_t)m * npix + i) * C + c]);
		if (c == 0 || acc > best) {
			best = acc;
			bi = c;
		}
	}
	labels[i] = (uint8_t)bi;
	if (masks != nullptr)
		for (int j = 0; j < n_mask; ++j) masks[(int64_t)j * npix + i] = (bi == mv.v[j]);
}

__global__ void __launch_bounds__(256) og_masks_kernel(
	const uint8_t *__restrict__ labels, int h, int w, int64_t pitch,
	uint8_t *__restrict__ masks, MaskValues mv, int n_mask) {
	const int64_t npix = (int64_t)h * w;
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= npix) return;
	const int y = (int)(i / w), x = (int)(i - (int64_t)y * w);
	const int v = labels[(int64_t)y * pitch + x];
	for (int j = 0; j < n_mask; ++j) masks[(int64_t)j * npix + i] = (v == mv.v[j]);
}

// Regions.binarized helpers (origami/batch/detect/layout.py:157-186)
__global__ void __launch_bounds__(256) og_label_ne_kernel(
	const uint8_t *__restrict__ labels, int h, int w, int64_t pitch, int bg, uint8_t *__restrict__ out, int64_t out_pitch) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
	if (x < w && y < h) out[(int64_t)y * out_pitch + x] = labels[(int64_t)y * pitch + x] != bg ? 255 : 0;
}

__global__ void __launch_bounds__(256) og_or_dilate3_kernel(
	const uint8_t *__restrict__ bin, int64_t bin_pitch, const uint8_t *__restrict__ mask, int64_t mask_pitch,
	int h, int w, float *__restrict__ out) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
	if (x >= w || y >= h) return;
	bool on = bin[(int64_t)y * bin_pitch + x] != 0;
	if (!on && mask != nullptr) {
		// skimage binary_dilation with square(3): 3x3 maximum, outside the image counts as 0
		for (int dy = -1; dy <= 1 && !on; ++dy) {
			const int yy = y + dy;
			if ((unsigned)yy >= (unsigned)h) continue;
			const uint8_t *r = mask + (int64_t)yy * mask_pitch;
			on = (x > 0 && r[x - 1]) || r[x] || (x + 1 < w && r[x + 1]);
		}
	}
	out[(int64_t)y * w + x] = on ? 1.f : 0.f;
}

struct AffineParams {
	double m00, m01, b1, m10, m11, b2;  // inverse map
};

__global__ void __launch_bounds__(256) og_warp_affine_kernel(
	const uint8_t *__restrict__ src, int H, int W, int64_t src_pitch, AffineParams a,
	uint8_t *__restrict__ dst, int width, int height, int64_t dst_pitch, int border) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x;
	const int y = blockIdx.y;
	if (x >= width || y >= height) return;
	// imgwarp.cpp warpAffine: AB_BITS = 10, round_delta = AB_SCALE / INTER_TAB_SIZE / 2 = 16
	const int adelta = __double2int_rn(__dmul_rn(__dmul_rn(a.m00, (double)x), 1024.0));
	const int bdelta = __double2int_rn(__dmul_rn(__dmul_rn(a.m10, (double)x), 1024.0));
	const int X0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m01, (double)y), a.b1), 1024.0)) + 16;
	const int Y0 = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.m11, (double)y), a.b2), 1024.0)) + 16;
	const int X = (X0 + adelta) >> 5, Y = (Y0 + bdelta) >> 5;
	og_fix q;
	q.ix = og_sat_short(X >> 5);
	q.iy = og_sat_short(Y >> 5);
	q.fx = X & 31;
	q.fy = Y & 31;
	const int p00 = og_tap(src, H, W, src_pitch, q.iy, q.ix, border);
	const int p01 = og_tap(src, H, W, src_pitch, q.iy, q.ix + 1, border);
	const int p10 = og_tap(src, H, W, src_pitch, q.iy + 1, q.ix, border);
	const int p11 = og_tap(src, H, W, src_pitch, q.iy + 1, q.ix + 1, border);
	dst[(int64_t)y * dst_pitch + x] = (uint8_t)og_blend(q, p00, p01, p10, p11);
}

int fill_masks(MaskValues &mv, const uint8_t *mask_values, int n_mask) {
	OG_REQUIRE(n_mask >= 0 && n_mask <= MAX_MASKS, OG_E_ARG, "n_mask %d outside [0, %d]", n_mask, MAX_MASKS);
	OG_REQUIRE(n_mask == 0 || mask_values != nullptr, OG_E_ARG, "mask_values is NULL");
	for (int j = 0; j < n_mask; ++j) mv.v[j] = mask_values[j];
	return OG_OK;
}

// Tile.write_inner (predict.py:60-63): argmax of a tile's inner rectangle straight into the page's
// label map.  One pixel per thread; consecutive threads walk a row of the rectangle.
__global__ void __launch_bounds__(256) og_argmax_tile_kernel(
	const float *__restrict__ probs, int M, int C, int th, int tw, int dx, int dy, int iw, int ih,
	uint8_t *__restrict__ labels, int64_t pitch, int x0, int y0) {
	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (int64_t)iw * ih) return;
	const int r = (int)(i / iw), c = (int)(i - (int64_t)r * iw);
	const int64_t tpix = (int64_t)th * tw, q = (int64_t)(dy + r) * tw + dx + c;
	int bi = 0;
	if (C == 4 && (reinterpret_cast<uintptr_t>(probs) & 15) == 0) {
		const float4 *p4 = reinterpret_cast<const float4 *>(probs);
		float4 acc = __ldcs(p4 + q);
		for (int m = 1; m < M; ++m) {
			const float4 v = __ldcs(p4 + (int64_t)m * tpix + q);
			acc.x = __fadd_rn(acc.x, v.x);
			acc.y = __fadd_rn(acc.y, v.y);
			acc.z = __fadd_rn(acc.z, v.z);
			acc.w = __fadd_rn(acc.w, v.w);
		}
		float best = acc.x;
		if (acc.y > best) { best = acc.y; bi = 1; }
		if (acc.z > best) { best = acc.z; bi = 2; }
		if (acc.w > best) { best = acc.w; bi = 3; }
	} else {
		float best = 0.f;
		for (int k = 0; k < C; ++k) {
			float acc = probs[q * C + k];
			for (int m = 1; m < M; ++m) acc = __fadd_rn(acc, probs[((int64_t)m * tpix + q) * C + k]);
			if (k == 0 || acc > best) {  // first maximum wins
				best = acc;
				bi = k;
			}
		}
	}
	labels[(int64_t)(y0 + r) * pitch + x0 + c] = (uint8_t)bi;
}

}  // namespace

extern "C" int og_argmax_tile(const float *probs, int M, int C, int th, int tw,
	int inner_dx, int inner_dy, int inner_w, int inner_h,
	uint8_t *labels, int lh, int lw, int64_t labels_pitch, int x0, int y0, void *stream) {
	OG_REQUIRE(probs != nullptr && labels != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(M >= 1 && C >= 1 && C <= 256 && th > 0 && tw > 0, OG_E_ARG, "tile shape (%d, %d, %d, %d) invalid", M, th, tw, C);
	OG_REQUIRE(inner_w >= 0 && inner_h >= 0 && inner_dx >= 0 && inner_dy >= 0 && inner_dx + inner_w <= tw && inner_dy + inner_h <= th,
		OG_E_ARG, "inner rectangle leaves the tile");
	OG_REQUIRE(lh > 0 && lw > 0 && labels_pitch >= lw && x0 >= 0 && y0 >= 0 && x0 + inner_w <= lw && y0 + inner_h <= lh,
		OG_E_ARG, "inner rectangle leaves the label map");
	const int64_t n = (int64_t)inner_w * inner_h;
	if (n == 0) return OG_OK;
	og_argmax_tile_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(probs, M, C, th, tw,
		inner_dx, inner_dy, inner_w, inner_h, labels, labels_pitch, x0, y0);
	return og_check_launch("og_argmax_tile_kernel");
}

extern "C" int og_argmax_labels(const float *probs, int M, int C, int h, int w,
	uint8_t *labels_out, uint8_t *masks_out, const uint8_t *mask_values, int n_mask, void *stream) {
	OG_REQUIRE(probs != nullptr && labels_out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(M >= 1 && C >= 1 && C <= 256 && h > 0 && w > 0, OG_E_ARG, "shape (%d, %d, %d, %d) invalid", M, h, w, C);
	MaskValues mv{};
	int rc = fill_masks(mv, mask_values, masks_out ? n_mask : 0);
	if (rc) return rc;
	const int64_t npix = (int64_t)h * w;
	cudaStream_t st = (cudaStream_t)stream;
	if (C == 4 && (reinterpret_cast<uintptr_t>(probs) & 15) == 0 && M <= 8) {
		const float4 *p4 = reinterpret_cast<const float4 *>(probs);
		const unsigned blocks = (unsigned)((npix + 1023) / 1024);
		switch (M) {
#define OG_ARGMAX_CASE(MT) case MT: og_argmax4_unrolled_kernel<MT><<<blocks, 256, 0, st>>>(p4, npix, labels_out, masks_out, mv, n_mask); break;
			OG_ARGMAX_CASE(1) OG_ARGMAX_CASE(2) OG_ARGMAX_CASE(3) OG_ARGMAX_CASE(4)
			OG_ARGMAX_CASE(5) OG_ARGMAX_CASE(6) OG_ARGMAX_CASE(7) OG_ARGMAX_CASE(8)
#undef OG_ARGMAX_CASE
		}
	} else if (C == 4 && (reinterpret_cast<uintptr_t>(probs) & 15) == 0) {
		const int64_t quads = (npix + 3) / 4;
		og_argmax4_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, st>>>(
			reinterpret_cast<const float4 *>(probs), M, npix, labels_out, masks_out, mv, n_mask);
	} else {
		og_argmax_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(probs, M, C, npix, labels_out, masks_out, mv, n_mask);
	}
	return og_check_launch("og_argmax_kernel");
}

extern "C" int og_class_masks(const uint8_t *labels, int h, int w, int64_t pitch,
	uint8_t *masks_out, const uint8_t *mask_values, int n_mask, void *stream) {
	OG_REQUIRE(labels != nullptr && masks_out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && pitch >= w, OG_E_ARG, "image %d x %d / pitch invalid", w, h);
	MaskValues mv{};
	int rc = fill_masks(mv, mask_values, n_mask);
	if (rc) return rc;
	if (n_mask == 0) return OG_OK;
	const int64_t npix = (int64_t)h * w;
	og_masks_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, (cudaStream_t)stream>>>(labels, h, w, pitch, masks_out, mv, n_mask);
	return og_check_launch("og_masks_kernel");
}

extern "C" int og_warp_affine(const uint8_t *src, int H, int W, int64_t src_pitch,
	const double *M, uint8_t *dst, int width, int height, int64_t dst_pitch, int border_value, void *stream) {
	OG_REQUIRE(src != nullptr && dst != nullptr && M != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(H > 0 && W > 0 && width > 0 && height > 0 && src_pitch >= W && dst_pitch >= width, OG_E_ARG, "size / pitch invalid");
	OG_REQUIRE(H < 32767 && W < 32767 && width < 32767 && height < 32767, OG_E_RANGE, "size exceeds OpenCV's 32767 limit");
	// invert the forward matrix exactly as imgwarp.cpp does (double precision)
	volatile double m0 = M[0], m1 = M[1], m2 = M[2], m3 = M[3], m4 = M[4], m5 = M[5];
	double D = m0 * m4 - m1 * m3;
	D = D != 0 ? 1. / D : 0;
	const double A11 = m4 * D, A22 = m0 * D;
	AffineParams a;
	a.m00 = A11;
	a.m01 = m1 * (-D);
	a.m10 = m3 * (-D);
	a.m11 = A22;
	volatile double t0 = -a.m00 * m2, t1 = a.m01 * m5, t2 = -a.m10 * m2, t3 = a.m11 * m5;
	a.b1 = t0 - t1;
	a.b2 = t2 - t3;
	dim3 block(256), grid((width + 255) / 256, height);
	og_warp_affine_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(src, H, W, src_pitch, a, dst, width, height, dst_pitch,
		border_value & 255);
	return og_check_launch("og_warp_affine_kernel");
}

extern "C" int og_label_mask_ne(const uint8_t *labels, int h, int w, int64_t pitch, int background,
	uint8_t *out, int64_t out_pitch, void *stream) {
	OG_REQUIRE(labels != nullptr && out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && pitch >= w && out_pitch >= w && h < 65536, OG_E_ARG, "image %d x %d / pitch invalid", w, h);
	dim3 block(256), grid((w + 255) / 256, h);
	og_label_ne_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(labels, h, w, pitch, background, out, out_pitch);
	return og_check_launch("og_label_ne_kernel");
}

extern "C" int og_or_dilate3(const uint8_t *bin, int64_t bin_pitch, const uint8_t *mask, int64_t mask_pitch,
	int h, int w, float *out, void *stream) {
	OG_REQUIRE(bin != nullptr && out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h > 0 && w > 0 && bin_pitch >= w && (mask == nullptr || mask_pitch >= w) && h < 65536, OG_E_ARG,
		"image %d x %d / pitch invalid", w, h);
	dim3 block(256), grid((w + 255) / 256, h);
	og_or_dilate3_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(bin, bin_pitch, mask, mask_pitch, h, w, out);
	return og_check_launch("og_or_dilate3_kernel");
}
