// Shared device arithmetic for the origami_b200 kernels (sm_100a).
//
// Everything here reproduces, operation for operation, the CPU arithmetic of the
// reference's page-geometry path; rounding-sensitive steps use
// the explicit *_rn intrinsics so that nvcc can never contract them into FMAs.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/origami_b200.h"

int og_set_error(int code, const char *fmt, ...);
int og_check_launch(const char *what);

#define OG_REQUIRE(cond, code, ...) \
	do { if (!(cond)) return og_set_error((code), __VA_ARGS__); } while (0)

// ---------------------------------------------------------------------------------------
// A2: cv2.resize INTER_LINEAR, generic path (resize.cpp).  Per destination index d:
// f = (float)((d + 0.5) * (1.0 / cell) - 0.5) in double, i = floor(f), f -= i in fp32.
// ---------------------------------------------------------------------------------------
struct og_axis {
	int i;    // source index (unclamped for the y axis)
	float f;  // fraction
};

__device__ __forceinline__ og_axis og_resize_axis(int d, double inv_cell) {
	double t = __dadd_rn(__dmul_rn(__dadd_rn((double)d, 0.5), inv_cell), -0.5);
	float f = __double2float_rn(t);
	float fl = floorf(f);
	og_axis a;
	a.i = (int)fl;
	a.f = __fsub_rn(f, fl);
	return a;
}

// x axis: index clamped and fraction reset at both borders (resize.cpp xofs/alpha setup).
__device__ __forceinline__ og_axis og_resize_axis_x(int d, double inv_cell, int gw) {
	og_axis a = og_resize_axis(d, inv_cell);
	if (a.i < 0) { a.i = 0; a.f = 0.f; }
	if (a.i >= gw - 1) { a.i = gw - 1; a.f = 0.f; }
	return a;
}

// fl(fl(a*(1-f)) + fl(b*f)) -- HResizeLinear / VResizeLinear without FMA.
__device__ __forceinline__ float og_lerp_cv(float a, float b, float w0, float w1) {
	return __fadd_rn(__fmul_rn(a, w0), __fmul_rn(b, w1));
}

// One node of the dense map, recomputed from the sparse grid (A2 for a single pixel).
__device__ __forceinline__ float2 og_dense_node(
	const float2 *__restrict__ grid, int gh, int gw, double inv_cell, int dy, int dx) {
	og_axis ax = og_resize_axis_x(dx, inv_cell, gw);
	og_axis ay = og_resize_axis(dy, inv_cell);
	int r0 = min(max(ay.i, 0), gh - 1);
	int r1 = min(max(ay.i + 1, 0), gh - 1);
	int c1 = min(ax.i + 1, gw - 1);
	float a0 = __fsub_rn(1.f, ax.f), a1 = ax.f;
	float b0 = __fsub_rn(1.f, ay.f), b1 = ay.f;
	float2 s00 = __ldg(grid + (size_t)r0 * gw + ax.i);
	float2 s01 = __ldg(grid + (size_t)r0 * gw + c1);
	float2 s10 = __ldg(grid + (size_t)r1 * gw + ax.i);
	float2 s11 = __ldg(grid + (size_t)r1 * gw + c1);
	float2 h0, h1, d;
	h0.x = og_lerp_cv(s00.x, s01.x, a0, a1);
	h0.y = og_lerp_cv(s00.y, s01.y, a0, a1);
	h1.x = og_lerp_cv(s10.x, s11.x, a0, a1);
	h1.y = og_lerp_cv(s10.y, s11.y, a0, a1);
	d.x = og_lerp_cv(h0.x, h1.x, b0, b1);
	d.y = og_lerp_cv(h0.y, h1.y, b0, b1);
	return d;
}

// ---------------------------------------------------------------------------------------
// A3: cv2.remap INTER_LINEAR on 8-bit (imgwarp.cpp RemapInvoker + remapBilinear).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ int og_sat_short(int v) {
	return min(max(v, -32768), 32767);
}

struct og_fix {
	int ix, iy, fx, fy;
};

__device__ __forceinline__ int og_blend(og_fix q, int p00, int p01, int p10, int p11) {
	int wx0 = 32 - q.fx, wy0 = 32 - q.fy;
	int acc = wy0 * (wx0 * p00 + q.fx * p01) + q.fy * (wx0 * p10 + q.fx * p11);
	return (acc + 512) >> 10;
}

__device__ __forceinline__ int og_tap(
	const uint8_t *__restrict__ src, int H, int W, int64_t pitch, int y, int x, int border) {
	if ((unsigned)x < (unsigned)W && (unsigned)y < (unsigned)H)
		return (int)__ldg(src + (int64_t)y * pitch + x);
	return border;
}

// N pixels at once: when every tap of every pixel is inside the image (the common case) all
// 4N byte loads are issued back to back before the first blend, so their latencies overlap;
// the address is formed in 32-bit arithmetic (H * pitch < 2^31 is checked on the host).  Border
// pixels take the per-tap test that yields OpenCV's BORDER_CONSTANT value per tap.
template <int N>
__device__ __forceinline__ void og_sample_linear_batch(
	const uint8_t *__restrict__ src, int H, int W, int pitch,
	const float (&mx)[N], const float (&my)[N], int (&out)[N]) {
	int ix[N], iy[N];
	og_fix q[N];
	bool all_in = true;
#pragma unroll
	for (int k = 0; k < N; ++k) {
		const int sx = __float2int_rn(__fmul_rn(mx[k], 32.f));
		const int sy = __float2int_rn(__fmul_rn(my[k], 32.f));
		ix[k] = sx >> 5;
		iy[k] = sy >> 5;
		q[k].fx = sx & 31;
		q[k].fy = sy & 31;
		all_in = all_in && (unsigned)ix[k] < (unsigned)(W - 1) && (unsigned)iy[k] < (unsigned)(H - 1);
	}
	if (all_in) {
		int p00[N], p01[N], p10[N], p11[N];
#pragma unroll
		for (int k = 0; k < N; ++k) {
			const uint8_t *s = src + (iy[k] * pitch + ix[k]);
			p00[k] = __ldg(s);
			p01[k] = __ldg(s + 1);
			p10[k] = __ldg(s + pitch);
			p11[k] = __ldg(s + pitch + 1);
		}
#pragma unroll
		for (int k = 0; k < N; ++k) out[k] = og_blend(q[k], p00[k], p01[k], p10[k], p11[k]);
	} else {
#pragma unroll
		for (int k = 0; k < N; ++k) {
			q[k].ix = og_sat_short(ix[k]);
			q[k].iy = og_sat_short(iy[k]);
			const int a = og_tap(src, H, W, pitch, q[k].iy, q[k].ix, 0);
			const int b = og_tap(src, H, W, pitch, q[k].iy, q[k].ix + 1, 0);
			const int c = og_tap(src, H, W, pitch, q[k].iy + 1, q[k].ix, 0);
			const int d = og_tap(src, H, W, pitch, q[k].iy + 1, q[k].ix + 1, 0);
			out[k] = og_blend(q[k], a, b, c, d);
		}
	}
}

// Split form of the sampler for software pipelining: load the taps now, blend them later.
struct og_taps {
	int p00, p01, p10, p11;
	int fx, fy;
};

// taps of N samples given in OpenCV's 1/32-px fixed point (sx = cvRound(x * 32))
template <int N>
__device__ __forceinline__ void og_load_taps_fixed(
	const uint8_t *__restrict__ src, int H, int W, int pitch,
	const int (&sx)[N], const int (&sy)[N], og_taps (&t)[N]) {
	int ix[N], iy[N];
	bool all_in = true;
#pragma unroll
	for (int k = 0; k < N; ++k) {
		ix[k] = sx[k] >> 5;
		iy[k] = sy[k] >> 5;
		t[k].fx = sx[k] & 31;
		t[k].fy = sy[k] & 31;
		all_in = all_in && (unsigned)ix[k] < (unsigned)(W - 1) && (unsigned)iy[k] < (unsigned)(H - 1);
	}
	if (all_in) {
#pragma unroll
		for (int k = 0; k < N; ++k) {
			const uint8_t *s = src + (iy[k] * pitch + ix[k]);
			t[k].p00 = __ldg(s);
			t[k].p01 = __ldg(s + 1);
			t[k].p10 = __ldg(s + pitch);
			t[k].p11 = __ldg(s + pitch + 1);
		}
	} else {
#pragma unroll
		for (int k = 0; k < N; ++k) {
			const int x = og_sat_short(ix[k]), y = og_sat_short(iy[k]);
			t[k].p00 = og_tap(src, H, W, pitch, y, x, 0);
			t[k].p01 = og_tap(src, H, W, pitch, y, x + 1, 0);
			t[k].p10 = og_tap(src, H, W, pitch, y + 1, x, 0);
			t[k].p11 = og_tap(src, H, W, pitch, y + 1, x + 1, 0);
		}
	}
}

template <int N>
__device__ __forceinline__ void og_load_taps_batch(
	const uint8_t *__restrict__ src, int H, int W, int pitch,
	const float (&mx)[N], const float (&my)[N], og_taps (&t)[N]) {
	int sx[N], sy[N];
#pragma unroll
	for (int k = 0; k < N; ++k) {
		sx[k] = __float2int_rn(__fmul_rn(mx[k], 32.f));
		sy[k] = __float2int_rn(__fmul_rn(my[k], 32.f));
	}
	og_load_taps_fixed<N>(src, H, W, pitch, sx, sy, t);
}

__device__ __forceinline__ int og_blend_taps(const og_taps &t) {
	const int wx0 = 32 - t.fx, wy0 = 32 - t.fy;
	const int acc = wy0 * (wx0 * t.p00 + t.fx * t.p01) + t.fy * (wx0 * t.p10 + t.fx * t.p11);
	return (acc + 512) >> 10;
}

// A4: cv2.remap INTER_NEAREST: saturate_cast<short>(float) = round half to even, OOB -> 0.
__device__ __forceinline__ int og_sample_nearest_fast(
	const uint8_t *__restrict__ src, int H, int W, int pitch, float mx, float my) {
	const int ix = og_sat_short(__float2int_rn(mx));
	const int iy = og_sat_short(__float2int_rn(my));
	if ((unsigned)ix < (unsigned)W && (unsigned)iy < (unsigned)H) return (int)__ldg(src + (iy * pitch + ix));
	return 0;
}

// ---------------------------------------------------------------------------------------
// A5: RegularGridInterpolator linear (scipy/interpolate/_rgi.py _evaluate_linear), float64,
// nodes recomputed from the sparse grid.  (y, x) dewarped -> (x, y) warped.
// ---------------------------------------------------------------------------------------
struct og_d2 {
	double x, y;
};

__device__ __forceinline__ og_d2 og_blend_f64(
	float2 v00, float2 v01, float2 v10, float2 v11, double a, double b) {
	double wa = __dsub_rn(1.0, a), wb = __dsub_rn(1.0, b);
	double w00 = __dmul_rn(wa, wb), w01 = __dmul_rn(wa, b);
	double w10 = __dmul_rn(a, wb), w11 = __dmul_rn(a, b);
	og_d2 o;
	o.x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00.x, w00), __dmul_rn((double)v01.x, w01)),
		__dmul_rn((double)v10.x, w10)), __dmul_rn((double)v11.x, w11));
	o.y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00.y, w00), __dmul_rn((double)v01.y, w01)),
		__dmul_rn((double)v10.y, w10)), __dmul_rn((double)v11.y, w11));
	return o;
}

__device__ __forceinline__ og_d2 og_inverse_point(
	const float2 *__restrict__ grid, int gh, int gw, int cell, double inv_cell, double y, double x) {
	int Hd = gh * cell, Wd = gw * cell;
	double fy = floor(y), fx = floor(x);
	fy = fmin(fmax(fy, 0.0), (double)(Hd - 2));
	fx = fmin(fmax(fx, 0.0), (double)(Wd - 2));
	int i0 = (int)fy, i1 = (int)fx;
	double a = __dsub_rn(y, fy), b = __dsub_rn(x, fx);
	float2 v00 = og_dense_node(grid, gh, gw, inv_cell, i0, i1);
	float2 v01 = og_dense_node(grid, gh, gw, inv_cell, i0, i1 + 1);
	float2 v10 = og_dense_node(grid, gh, gw, inv_cell, i0 + 1, i1);
	float2 v11 = og_dense_node(grid, gh, gw, inv_cell, i0 + 1, i1 + 1);
	return og_blend_f64(v00, v01, v10, v11, a, b);
}

// ---------------------------------------------------------------------------------------
// A8: np.linspace([0, 0], stop, num) (numpy/_core/function_base.py) and the lattice of
// Line.warped_grid (origami/core/block.py:238-241).
// ---------------------------------------------------------------------------------------
struct og_lin {
	double stepx, stepy;  // stop / div
	double stopx, stopy;
	double div;
	int num;
	int any_zero;  // any step component == 0 -> (i/div)*stop branch for every component
};

__device__ __forceinline__ og_lin og_linspace_setup(double sx, double sy, int num) {
	og_lin l;
	l.num = num;
	l.stopx = sx;
	l.stopy = sy;
	l.div = (double)(num - 1);
	if (num > 1) {
		l.stepx = __ddiv_rn(sx, l.div);
		l.stepy = __ddiv_rn(sy, l.div);
		l.any_zero = (l.stepx == 0.0) || (l.stepy == 0.0);
	} else {
		l.stepx = sx;  // div == 0: y = i * delta with i = 0
		l.stepy = sy;
		l.any_zero = 0;
	}
	return l;
}

__device__ __forceinline__ og_d2 og_linspace_at(const og_lin &l, int i) {
	og_d2 o;
	if (l.num > 1 && i == l.num - 1) {
		o.x = l.stopx;
		o.y = l.stopy;
		return o;
	}
	double di = (double)i;
	if (l.any_zero) {
		double t = __ddiv_rn(di, l.div);
		o.x = __dmul_rn(t, l.stopx);
		o.y = __dmul_rn(t, l.stopy);
	} else {
		o.x = __dmul_rn(di, l.stepx);
		o.y = __dmul_rn(di, l.stepy);
	}
	// `y += start` with start = 0.0 turns -0.0 into +0.0; harmless but kept.
	o.x = __dadd_rn(o.x, 0.0);
	o.y = __dadd_rn(o.y, 0.0);
	return o;
}
