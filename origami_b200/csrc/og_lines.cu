// Line kernels: step counts (A8/A21), warped sampling grids (A8), strip extraction
// (A9/A10), label sampling (A15) and the bare inverse lookup (A5).
//
// Reference: Line.warped_grid / Line.dewarped_image origami/core/block.py:223-259,
// Grid.inverse_yx origami/core/dewarp.py:561-567, LineExtractor origami/batch/core/
// lines.py:131-149, ConfidenceSampler origami/batch/detect/lines.py:34-54.
//
// The reference materialises the 281 MB dense map and feeds it to SciPy's interpolator for
// every line.  Here a work item is (line, 64-column chunk): the CTA recomputes, with the
// exact cv2.resize arithmetic, only the dense-map nodes under that chunk into shared memory
// (as float64, so each fp32->fp64 conversion happens once per node instead of four times
// per sample), then every thread walks a column of the strip doing SciPy's float64 blend,
// the fp32 cast, OpenCV's 1/32-px quantisation and the integer bilinear blend.  Work items
// come from a device-side prefix sum, so a whole page is ONE launch with no host round trip.
#include "og_common.cuh"

namespace {

constexpr int CW = 64;          // strip columns per work item
constexpr int LTHREADS = 256;   // threads per CTA
constexpr int RSPLIT = LTHREADS / CW;
constexpr int SMEM_BUDGET = 100 * 1024;
constexpr int MAX_CLASSES = 256;

// ---------------------------------------------------------------------------------------
// step counts + prefix sums: one CTA
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float2 rough_point(const float2 *grid, int gh, int gw, int cell,
	double inv_cell, double px, double py) {
	og_d2 o = og_inverse_point(grid, gh, gw, cell, inv_cell, py, px);
	return make_float2(__double2float_rn(o.x), __double2float_rn(o.y));
}

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T *warp_sums, T &total) {
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	T inc = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		T t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) warp_sums[wid] = inc;
	__syncthreads();
	if (wid == 0) {
		T w = (lane < (blockDim.x >> 5)) ? warp_sums[lane] : T(0);
		T winc = w;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			T t = __shfl_up_sync(0xffffffffu, winc, o);
			if (lane >= o) winc += t;
		}
		warp_sums[lane] = winc - w;  // exclusive
		if (lane == 31) warp_sums[32] = winc;
	}
	__syncthreads();
	T res = warp_sums[wid] + inc - v;
	total = warp_sums[32];
	__syncthreads();
	return res;
}

__global__ void __launch_bounds__(1024) og_line_steps_kernel(
	const float2 *__restrict__ grid, int gh, int gw, int cell,
	const og_line *__restrict__ lines, int n,
	int32_t *__restrict__ steps_out, int64_t *__restrict__ pix_offsets, int32_t *__restrict__ chunk_offsets) {
	__shared__ long long ws64[33];
	__shared__ int ws32[33];
	const double inv_cell = 1.0 / (double)cell;
	long long carry_pix = 0;
	int carry_chunks = 0;
	for (int base = 0; base < n; base += blockDim.x) {
		const int i = base + threadIdx.x;
		int xs = 0, ys = 0;
		if (i < n) {
			const og_line ln = lines[i];
			xs = ln.xsteps;
			ys = ln.ysteps;
			if (xs <= 0 || ys <= 0) {
				// Line.warped_grid(xsteps=2, ysteps=2) on the un-clipped line (block.py:226-228):
				// samples p, p+right, (up+p), (up+p)+right
				const double qx = __dadd_rn(ln.up[0], ln.p[0]), qy = __dadd_rn(ln.up[1], ln.p[1]);
				float2 g00 = rough_point(grid, gh, gw, cell, inv_cell, ln.p[0], ln.p[1]);
				float2 g01 = rough_point(grid, gh, gw, cell, inv_cell, __dadd_rn(ln.p[0], ln.right[0]), __dadd_rn(ln.p[1], ln.right[1]));
				float2 g10 = rough_point(grid, gh, gw, cell, inv_cell, qx, qy);
				float2 g11 = rough_point(grid, gh, gw, cell, inv_cell, __dadd_rn(qx, ln.right[0]), __dadd_rn(qy, ln.right[1]));
				if (xs <= 0) {
					float wx = fmaxf(fabsf(__fsub_rn(g00.x, g01.x)), fabsf(__fsub_rn(g10.x, g11.x)));
					float v = ceilf(__fmul_rn(wx, (float)ln.xres));
					xs = (v >= 0.f && v < 1.0e6f) ? max(2, (int)v) : 0;
					if (!(wx == wx)) xs = 0;
				}
				if (ys <= 0) {
					float wy = fmaxf(fabsf(__fsub_rn(g00.y, g10.y)), fabsf(__fsub_rn(g01.y, g11.y)));
					float v = ceilf(__fmul_rn(wy, (float)ln.yres));
					ys = (v >= 0.f && v < 1.0e6f) ? max(2, (int)v) : 0;
					if (!(wy == wy)) ys = 0;
				}
			}
			steps_out[2 * i] = xs;
			steps_out[2 * i + 1] = ys;
		}
		long long npix = (long long)xs * (long long)ys;
		int nchunks = (xs > 0 && ys > 0) ? (xs + CW - 1) / CW : 0;
		long long tot_pix;
		int tot_chunks;
		long long ep = block_exclusive_scan<long long>(npix, ws64, tot_pix);
		int ec = block_exclusive_scan<int>(nchunks, ws32, tot_chunks);
		if (i < n) {
			pix_offsets[i] = carry_pix + ep;
			chunk_offsets[i] = carry_chunks + ec;
		}
		carry_pix += tot_pix;
		carry_chunks += tot_chunks;
	}
	if (threadIdx.x == 0) {
		pix_offsets[n] = carry_pix;
		chunk_offsets[n] = carry_chunks;
	}
}

// ---------------------------------------------------------------------------------------
// the per-(line, chunk) engine
// ---------------------------------------------------------------------------------------
enum { MODE_GRID = 0, MODE_STRIP = 1, MODE_LABELS = 2 };

struct LineParams {
	const float2 *grid;
	int gh, gw, cell;
	const og_line *lines;
	int n;
	const int32_t *steps;
	const int64_t *pix_offsets;
	const int32_t *chunk_offsets;
	// MODE_GRID
	float *grids_out;
	// MODE_STRIP
	const uint8_t *src;
	int H, W;
	int64_t src_pitch;
	uint8_t *strips_out;
	int32_t *bbox_out;
	int64_t capacity;
	// MODE_LABELS
	float scale_x, scale_y;
	int n_classes;
	unsigned long long *counts_out;
	unsigned long long *overflow_out;
};

struct RowEntry {
	int r0, r1;  // sparse rows relative to the staged band
	float b0, b1;
};

__device__ __forceinline__ int find_line(const int32_t *__restrict__ chunk_offsets, int n, int item) {
	int lo = 0, hi = n;  // invariant: chunk_offsets[lo] <= item < chunk_offsets[hi]
	while (hi - lo > 1) {
		int mid = (lo + hi) >> 1;
		if (__ldg(chunk_offsets + mid) <= item) lo = mid; else hi = mid;
	}
	return lo;
}

template <int MODE>
__global__ void __launch_bounds__(LTHREADS) og_lines_kernel(LineParams p) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	__shared__ unsigned int hist[MODE == MODE_LABELS ? MAX_CLASSES + 1 : 1];
	__shared__ int s_bbox[4];

	const int Hd = p.gh * p.cell, Wd = p.gw * p.cell;
	const double inv_cell = 1.0 / (double)p.cell;
	const int total_items = __ldg(p.chunk_offsets + p.n);
	const int tid = threadIdx.x;

	for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
		const int li = find_line(p.chunk_offsets, p.n, item);
		const int chunk = item - __ldg(p.chunk_offsets + li);
		const int xsteps = __ldg(p.steps + 2 * li), ysteps = __ldg(p.steps + 2 * li + 1);
		const int64_t pix0 = (MODE != MODE_LABELS) ? __ldg(p.pix_offsets + li) : 0;
		if (MODE != MODE_LABELS && pix0 + (int64_t)xsteps * ysteps > p.capacity) continue;
		const og_line ln = p.lines[li];
		const int c0 = chunk * CW;
		const int ncols = min(CW, xsteps - c0);
		const og_lin lx = og_linspace_setup(ln.cright[0], ln.cright[1], xsteps);
		const og_lin ly = og_linspace_setup(ln.up[0], ln.up[1], ysteps);

		// ---- region of dense-map nodes under this chunk (corners of the affine lattice)
		double ymin = 1e300, ymax = -1e300, xmin = 1e300, xmax = -1e300;
#pragma unroll
		for (int k = 0; k < 4; ++k) {
			og_d2 yr = og_linspace_at(ly, (k & 1) ? ysteps - 1 : 0);
			og_d2 xc = og_linspace_at(lx, (k & 2) ? c0 + ncols - 1 : c0);
			double px = __dadd_rn(__dadd_rn(yr.x, ln.cp[0]), xc.x);
			double py = __dadd_rn(__dadd_rn(yr.y, ln.cp[1]), xc.y);
			ymin = fmin(ymin, py); ymax = fmax(ymax, py);
			xmin = fmin(xmin, px); xmax = fmax(xmax, px);
		}
		bool staged = (ymin == ymin) && (xmin == xmin) && fabs(ymin) < 1e9 && fabs(ymax) < 1e9 &&
			fabs(xmin) < 1e9 && fabs(xmax) < 1e9;
		int R0 = 0, C0 = 0, NR = 0, NC = 0, SR0 = 0, NSR = 0;
		if (staged) {
			R0 = min(max((int)floor(ymin) - 1, 0), Hd - 2);
			int R1 = min(max((int)floor(ymax) + 1, 0), Hd - 2) + 1;
			C0 = min(max((int)floor(xmin) - 1, 0), Wd - 2);
			int C1 = min(max((int)floor(xmax) + 1, 0), Wd - 2) + 1;
			NR = R1 - R0 + 1;
			NC = C1 - C0 + 1;
			og_axis a0 = og_resize_axis(R0, inv_cell), a1 = og_resize_axis(R1, inv_cell);
			SR0 = min(max(a0.i, 0), p.gh - 1);
			int SR1 = min(max(a1.i + 1, 0), p.gh - 1);
			NSR = SR1 - SR0 + 1;
			size_t need = (size_t)NR * NC * sizeof(double2) + (size_t)NSR * NC * sizeof(float2) +
				(size_t)NR * sizeof(RowEntry);
			staged = need <= (size_t)SMEM_BUDGET;
		}
		double2 *nodes = reinterpret_cast<double2 *>(smem_raw);
		float2 *hs = reinterpret_cast<float2 *>(smem_raw + (size_t)NR * NC * sizeof(double2));
		RowEntry *rows = reinterpret_cast<RowEntry *>(smem_raw + (size_t)NR * NC * sizeof(double2) +
			(size_t)NSR * NC * sizeof(float2));

		__syncthreads();  // previous item's shared data is dead
		if (MODE == MODE_LABELS) {
			for (int k = tid; k <= MAX_CLASSES; k += LTHREADS) hist[k] = 0;
		}
		if (MODE == MODE_STRIP && tid < 4) s_bbox[tid] = (tid < 2) ? INT_MAX : INT_MIN;
		if (staged) {
			// horizontal pass over the sparse rows of the band (HResizeLinear)
			for (int j = tid; j < NC; j += LTHREADS) {
				og_axis ax = og_resize_axis_x(C0 + j, inv_cell, p.gw);
				const int c1 = min(ax.i + 1, p.gw - 1);
				const float w0 = __fsub_rn(1.f, ax.f), w1 = ax.f;
				for (int r = 0; r < NSR; ++r) {
					const float2 *g = p.grid + (size_t)(SR0 + r) * p.gw;
					float2 s0 = __ldg(g + ax.i), s1 = __ldg(g + c1);
					hs[r * NC + j] = make_float2(og_lerp_cv(s0.x, s1.x, w0, w1), og_lerp_cv(s0.y, s1.y, w0, w1));
				}
			}
			for (int i = tid; i < NR; i += LTHREADS) {
				og_axis ay = og_resize_axis(R0 + i, inv_cell);
				RowEntry e;
				e.r0 = min(max(ay.i, 0), p.gh - 1) - SR0;
				e.r1 = min(max(ay.i + 1, 0), p.gh - 1) - SR0;
				e.b1 = ay.f;
				e.b0 = __fsub_rn(1.f, ay.f);
				rows[i] = e;
			}
			__syncthreads();
			// vertical pass (VResizeLinear) -> float64 nodes
			for (int k = tid; k < NR * NC; k += LTHREADS) {
				const int i = k / NC, j = k - i * NC;
				const RowEntry e = rows[i];
				const float2 u = hs[e.r0 * NC + j], v = hs[e.r1 * NC + j];
				nodes[k] = make_double2((double)og_lerp_cv(u.x, v.x, e.b0, e.b1),
					(double)og_lerp_cv(u.y, v.y, e.b0, e.b1));
			}
		}
		__syncthreads();

		// ---- sample phase: thread = (column, row quarter)
		const int col = tid % CW, rq = tid / CW;
		const int rows_per = (ysteps + RSPLIT - 1) / RSPLIT;
		const int rbeg = rq * rows_per, rend = min(ysteps, rbeg + rows_per);
		float bx0 = 3.0e38f, by0 = 3.0e38f, bx1 = -3.0e38f, by1 = -3.0e38f;
		if (col < ncols) {
			const int c = c0 + col;
			const og_d2 xc = og_linspace_at(lx, c);
			for (int r = rbeg; r < rend; ++r) {
				const og_d2 yr = og_linspace_at(ly, r);
				// dewarped_grid = (ys + p0)[:, None] + xs[None, :]   (block.py:241)
				const double px = __dadd_rn(__dadd_rn(yr.x, ln.cp[0]), xc.x);
				const double py = __dadd_rn(__dadd_rn(yr.y, ln.cp[1]), xc.y);
				og_d2 w;
				bool done = false;
				if (staged) {
					double fy = fmin(fmax(floor(py), 0.0), (double)(Hd - 2));
					double fx = fmin(fmax(floor(px), 0.0), (double)(Wd - 2));
					const int ri = (int)fy - R0, ci = (int)fx - C0;
					if (ri >= 0 && ri < NR - 1 && ci >= 0 && ci < NC - 1) {
						const double a = __dsub_rn(py, fy), b = __dsub_rn(px, fx);
						const double wa = __dsub_rn(1.0, a), wb = __dsub_rn(1.0, b);
						const double w00 = __dmul_rn(wa, wb), w01 = __dmul_rn(wa, b);
						const double w10 = __dmul_rn(a, wb), w11 = __dmul_rn(a, b);
						const double2 v00 = nodes[ri * NC + ci], v01 = nodes[ri * NC + ci + 1];
						const double2 v10 = nodes[(ri + 1) * NC + ci], v11 = nodes[(ri + 1) * NC + ci + 1];
						w.x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(v00.x, w00), __dmul_rn(v01.x, w01)),
							__dmul_rn(v10.x, w10)), __dmul_rn(v11.x, w11));
						w.y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(v00.y, w00), __dmul_rn(v01.y, w01)),
							__dmul_rn(v10.y, w10)), __dmul_rn(v11.y, w11));
						done = true;
					}
				}
				if (!done) w = og_inverse_point(p.grid, p.gh, p.gw, p.cell, inv_cell, py, px);
				const float gx = __double2float_rn(w.x), gy = __double2float_rn(w.y);
				if (MODE == MODE_GRID) {
					float2 *o = reinterpret_cast<float2 *>(p.grids_out) + pix0 + (int64_t)r * xsteps + c;
					*o = make_float2(gx, gy);
				} else if (MODE == MODE_STRIP) {
					const int v = og_sample_linear(p.src, p.H, p.W, p.src_pitch, gx, gy, 0);
					// pixels[::-1, :]  (block.py:257)
					p.strips_out[pix0 + (int64_t)(ysteps - 1 - r) * xsteps + c] = (uint8_t)v;
					bx0 = fminf(bx0, gx); bx1 = fmaxf(bx1, gx);
					by0 = fminf(by0, gy); by1 = fmaxf(by1, gy);
				} else {
					// scale_grid (batch/detect/lines.py:18-22) then cv2.remap INTER_NEAREST
					const float sx = __fmul_rn(gx, p.scale_x), sy = __fmul_rn(gy, p.scale_y);
					const int lab = og_sample_nearest(p.src, p.H, p.W, p.src_pitch, sx, sy, 0);
					atomicAdd(&hist[min(lab, p.n_classes)], 1u);
				}
			}
		}
		if (MODE == MODE_STRIP && p.bbox_out != nullptr) {
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				bx0 = fminf(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
				by0 = fminf(by0, __shfl_xor_sync(0xffffffffu, by0, o));
				bx1 = fmaxf(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
				by1 = fmaxf(by1, __shfl_xor_sync(0xffffffffu, by1, o));
			}
			if ((tid & 31) == 0 && bx0 <= bx1) {
				atomicMin(&s_bbox[0], (int)floorf(bx0));
				atomicMin(&s_bbox[1], (int)floorf(by0));
				atomicMax(&s_bbox[2], (int)ceilf(bx1));
				atomicMax(&s_bbox[3], (int)ceilf(by1));
			}
			__syncthreads();
			if (tid < 4 && s_bbox[0] <= s_bbox[2]) {
				if (tid < 2) atomicMin(p.bbox_out + 4 * li + tid, s_bbox[tid]);
				else atomicMax(p.bbox_out + 4 * li + tid, s_bbox[tid]);
			}
		}
		if (MODE == MODE_LABELS) {
			__syncthreads();
			for (int k = tid; k < p.n_classes; k += LTHREADS)
				if (hist[k]) atomicAdd(p.counts_out + (size_t)li * p.n_classes + k, (unsigned long long)hist[k]);
			if (tid == 0 && hist[p.n_classes] && p.overflow_out != nullptr)
				atomicAdd(p.overflow_out + li, (unsigned long long)hist[p.n_classes]);
		}
	}
}

__global__ void og_bbox_init_kernel(int32_t *bbox, int n) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) {
		bbox[4 * i + 0] = INT_MAX;
		bbox[4 * i + 1] = INT_MAX;
		bbox[4 * i + 2] = INT_MIN;
		bbox[4 * i + 3] = INT_MIN;
	}
}

__global__ void og_inverse_kernel(const float2 *__restrict__ grid, int gh, int gw, int cell,
	const double *__restrict__ pts, int64_t n, double *__restrict__ out) {
	int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const double inv_cell = 1.0 / (double)cell;
	og_d2 o = og_inverse_point(grid, gh, gw, cell, inv_cell, pts[2 * i], pts[2 * i + 1]);
	out[2 * i] = o.x;
	out[2 * i + 1] = o.y;
}

int check_grid(const float *grid, int gh, int gw, int cell) {
	OG_REQUIRE(grid != nullptr, OG_E_ARG, "grid is NULL");
	OG_REQUIRE(gh >= 2 && gw >= 2 && cell >= 1, OG_E_ARG, "grid shape (%d, %d) cell %d invalid", gh, gw, cell);
	OG_REQUIRE((reinterpret_cast<uintptr_t>(grid) & 7) == 0, OG_E_ALIGN, "grid must be 8-byte aligned");
	OG_REQUIRE((int64_t)gh * cell < 32767 && (int64_t)gw * cell < 32767, OG_E_RANGE,
		"dewarped size exceeds OpenCV's 32767 limit");
	return OG_OK;
}

int check_lines(const og_line *lines, int n, const int32_t *steps, const int32_t *chunk_offsets) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	OG_REQUIRE(n == 0 || lines != nullptr, OG_E_ARG, "lines is NULL");
	OG_REQUIRE(steps != nullptr && chunk_offsets != nullptr, OG_E_ARG, "steps / chunk_offsets is NULL");
	OG_REQUIRE((reinterpret_cast<uintptr_t>(lines) & 7) == 0, OG_E_ALIGN, "lines must be 8-byte aligned");
	return OG_OK;
}

template <int MODE>
int launch_lines(const LineParams &p, cudaStream_t st) {
	cudaError_t e = cudaFuncSetAttribute(og_lines_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
	if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
	// persistent grid: two CTAs per SM fit the 100 KB staging budget
	og_lines_kernel<MODE><<<OG_NUM_SMS * 2, LTHREADS, SMEM_BUDGET, st>>>(p);
	return og_check_launch("og_lines_kernel");
}

}  // namespace

extern "C" int og_inverse_yx(const float *grid, int gh, int gw, int cell,
	const double *pts_yx, int64_t n, double *out_xy, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	if (n == 0) return OG_OK;
	OG_REQUIRE(pts_yx != nullptr && out_xy != nullptr, OG_E_ARG, "pts_yx / out_xy is NULL");
	const int threads = 256;
	og_inverse_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
		reinterpret_cast<const float2 *>(grid), gh, gw, cell, pts_yx, n, out_xy);
	return og_check_launch("og_inverse_kernel");
}

extern "C" int og_line_steps(const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n,
	int32_t *steps_out, int64_t *pix_offsets_out, int32_t *chunk_offsets_out, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	OG_REQUIRE(n == 0 || lines != nullptr, OG_E_ARG, "lines is NULL");
	OG_REQUIRE(pix_offsets_out != nullptr && chunk_offsets_out != nullptr && (n == 0 || steps_out != nullptr),
		OG_E_ARG, "output pointer is NULL");
	og_line_steps_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2 *>(grid), gh, gw, cell,
		lines, n, steps_out, pix_offsets_out, chunk_offsets_out);
	return og_check_launch("og_line_steps_kernel");
}

extern "C" int og_warped_grids(const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, const int32_t *chunk_offsets,
	float *grids_out, int64_t capacity, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	if (n == 0) return OG_OK;
	OG_REQUIRE(pix_offsets != nullptr && grids_out != nullptr, OG_E_ARG, "pix_offsets / grids_out is NULL");
	OG_REQUIRE((reinterpret_cast<uintptr_t>(grids_out) & 7) == 0, OG_E_ALIGN, "grids_out must be 8-byte aligned");
	LineParams p{};
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh; p.gw = gw; p.cell = cell;
	p.lines = lines; p.n = n;
	p.steps = steps; p.pix_offsets = pix_offsets; p.chunk_offsets = chunk_offsets;
	p.grids_out = grids_out;
	p.capacity = capacity;
	return launch_lines<MODE_GRID>(p, (cudaStream_t)stream);
}

extern "C" int og_extract_lines(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, const int32_t *chunk_offsets,
	uint8_t *strips_out, int64_t capacity, int32_t *bbox_out, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	if (n == 0) return OG_OK;
	OG_REQUIRE(src != nullptr && H > 0 && W > 0 && src_pitch >= W, OG_E_ARG, "src invalid");
	OG_REQUIRE(H < 32767 && W < 32767, OG_E_RANGE, "source %d x %d exceeds OpenCV's 32767 limit", W, H);
	OG_REQUIRE(pix_offsets != nullptr && strips_out != nullptr, OG_E_ARG, "pix_offsets / strips_out is NULL");
	cudaStream_t st = (cudaStream_t)stream;
	if (bbox_out != nullptr) {
		og_bbox_init_kernel<<<(n + 255) / 256, 256, 0, st>>>(bbox_out, n);
		rc = og_check_launch("og_bbox_init_kernel");
		if (rc) return rc;
	}
	LineParams p{};
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh; p.gw = gw; p.cell = cell;
	p.lines = lines; p.n = n;
	p.steps = steps; p.pix_offsets = pix_offsets; p.chunk_offsets = chunk_offsets;
	p.src = src; p.H = H; p.W = W; p.src_pitch = src_pitch;
	p.strips_out = strips_out;
	p.bbox_out = bbox_out;
	p.capacity = capacity;
	return launch_lines<MODE_STRIP>(p, st);
}

extern "C" int og_sample_labels(const uint8_t *labels, int lh, int lw, int64_t labels_pitch,
	int H, int W,
	const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n,
	const int32_t *steps, const int32_t *chunk_offsets,
	int n_classes, int64_t *counts_out, int64_t *overflow_out, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	OG_REQUIRE(n_classes >= 1 && n_classes <= MAX_CLASSES, OG_E_ARG, "n_classes %d outside [1, %d]", n_classes, MAX_CLASSES);
	if (n == 0) return OG_OK;
	OG_REQUIRE(labels != nullptr && lh > 0 && lw > 0 && labels_pitch >= lw, OG_E_ARG, "labels invalid");
	OG_REQUIRE(lh < 32767 && lw < 32767, OG_E_RANGE, "label map exceeds OpenCV's 32767 limit");
	OG_REQUIRE(H > 0 && W > 0, OG_E_ARG, "page size invalid");
	OG_REQUIRE(counts_out != nullptr, OG_E_ARG, "counts_out is NULL");
	cudaStream_t st = (cudaStream_t)stream;
	cudaError_t e = cudaMemsetAsync(counts_out, 0, sizeof(int64_t) * (size_t)n * n_classes, st);
	if (e == cudaSuccess && overflow_out != nullptr) e = cudaMemsetAsync(overflow_out, 0, sizeof(int64_t) * (size_t)n, st);
	if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
	LineParams p{};
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh; p.gw = gw; p.cell = cell;
	p.lines = lines; p.n = n;
	p.steps = steps; p.chunk_offsets = chunk_offsets;
	p.pix_offsets = nullptr;
	p.src = labels; p.H = lh; p.W = lw; p.src_pitch = labels_pitch;
	// scale_grid: grid[:, :, 0] *= w1 / w0 with the ratio rounded to fp32
	p.scale_x = (float)((double)lw / (double)W);
	p.scale_y = (float)((double)lh / (double)H);
	p.n_classes = n_classes;
	p.counts_out = reinterpret_cast<unsigned long long *>(counts_out);
	p.overflow_out = reinterpret_cast<unsigned long long *>(overflow_out);
	return launch_lines<MODE_LABELS>(p, st);
}
