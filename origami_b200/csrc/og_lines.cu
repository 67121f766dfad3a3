// Line kernels: step counts (A8/A21), warped sampling grids (A8), strip extraction
// (A9/A10), label sampling (A15) and the bare inverse lookup (A5).
//
// Reference: Line.warped_grid / Line.dewarped_image origami/core/block.py:223-259,
// Grid.inverse_yx origami/core/dewarp.py:561-567, LineExtractor origami/batch/core/
// lines.py:131-149, ConfidenceSampler origami/batch/detect/lines.py:34-54.
//
// The reference materialises the 281 MB dense map and feeds it to SciPy's interpolator for
// every line.  Here a work item is (line, 64-column chunk).  A persistent CTA of 8 worker
// warps + 1 planner warp loops over items:
//   planner  -- while the workers process item k, computes the descriptor of item k+1
//               (line lookup, np.linspace parameters, per-row lattice offsets, the region of
//               dense-map nodes under the chunk) into a double-buffered shared structure;
//   workers  -- recompute, with the exact cv2.resize arithmetic, only the dense-map nodes
//               under the chunk into shared memory (as float64, so each fp32->fp64
//               conversion happens once per node instead of four times per sample), then
//               every thread walks a column of the strip: SciPy's float64 blend, the fp32
//               cast, OpenCV's 1/32-px quantisation and the integer bilinear blend.
// Work items come from a device-side prefix sum, so a whole page is ONE launch with no host
// round trip.  The bound of this kernel is the FP64 pipe + instruction issue, not HBM: the
// float64 blend that bit-parity with SciPy requires costs ~30 DP-pipe instructions per
// sample.
//
// That is the engine for callers without a dense map (og_lines_kernel).  With the map
// (the page kernel emits it) the nodes are gathered from it: og_lines_map_kernel evaluates
// every sample in float64 (warped grids, strips with bounding boxes); strips and label
// counts go through og_strips_screened_kernel, which decides cv2.remap's fixed-point
// coordinate from a float32 interval and keeps the float64 sequence for the few samples the
// interval cannot decide -- same bits, two thirds of the time.
#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

constexpr int CW = 64;                  // strip columns per work item
constexpr int WORKERS = 256;            // worker threads per CTA
constexpr int LTHREADS = WORKERS + 32;  // + one planner warp
constexpr int RSPLIT = WORKERS / CW;    // rows are dealt round-robin to 4 thread groups
constexpr int SMEM_BUDGET = 100 * 1024;
constexpr int MAX_ROWS = 128;           // rows with a precomputed lattice offset
constexpr int MAX_CLASSES = 256;
#ifndef OG_RB
#define OG_RB 2
#endif
constexpr int RB = OG_RB;                   // strip rows a thread processes together (load batching)

__device__ __forceinline__ void bar_sync(int id, int nthreads) {
	asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------
// step counts: 4 lanes per line (one per corner of the rough 2x2 grid), then one scan CTA
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void og_line_steps_body(
	const float2 *__restrict__ grid, int gh, int gw, int cell,
	const og_line *__restrict__ lines, int n, int32_t *steps_out, int t) {
	const int i = min(t >> 2, n - 1), corner = t & 3;
	const double inv_cell = 1.0 / (double)cell;
	const og_line *ln = lines + i;
	int xs = ln->xsteps, ys = ln->ysteps;
	const bool derive = xs <= 0 || ys <= 0;
	// Line.warped_grid(xsteps=2, ysteps=2) on the un-clipped line (block.py:226-228):
	// samples p, p + right, (up + p), (up + p) + right
	double qx = ln->p[0], qy = ln->p[1];
	if (corner & 2) {
		qx = __dadd_rn(ln->up[0], qx);
		qy = __dadd_rn(ln->up[1], qy);
	}
	if (corner & 1) {
		qx = __dadd_rn(qx, ln->right[0]);
		qy = __dadd_rn(qy, ln->right[1]);
	}
	float gx = 0.f, gy = 0.f;
	if (derive) {
		og_d2 o = og_inverse_point(grid, gh, gw, cell, inv_cell, qy, qx);
		gx = __double2float_rn(o.x);
		gy = __double2float_rn(o.y);
	}
	const int base = (threadIdx.x & 31) & ~3;
	const float g00x = __shfl_sync(0xffffffffu, gx, base), g00y = __shfl_sync(0xffffffffu, gy, base);
	const float g01x = __shfl_sync(0xffffffffu, gx, base + 1), g01y = __shfl_sync(0xffffffffu, gy, base + 1);
	const float g10x = __shfl_sync(0xffffffffu, gx, base + 2), g10y = __shfl_sync(0xffffffffu, gy, base + 2);
	const float g11x = __shfl_sync(0xffffffffu, gx, base + 3), g11y = __shfl_sync(0xffffffffu, gy, base + 3);
	if ((t >> 2) < n && corner == 0) {
		if (xs <= 0) {
			// np.max(np.abs(rough[:, 0, 0] - rough[:, 1, 0])) * xres, all fp32
			const float wx = fmaxf(fabsf(__fsub_rn(g00x, g01x)), fabsf(__fsub_rn(g10x, g11x)));
			const float v = ceilf(__fmul_rn(wx, (float)ln->xres));
			xs = (wx == wx && v >= 0.f && v < 1.0e6f) ? max(2, (int)v) : 0;
		}
		if (ys <= 0) {
			const float wy = fmaxf(fabsf(__fsub_rn(g00y, g10y)), fabsf(__fsub_rn(g01y, g11y)));
			const float v = ceilf(__fmul_rn(wy, (float)ln->yres));
			ys = (wy == wy && v >= 0.f && v < 1.0e6f) ? max(2, (int)v) : 0;
		}
		if ((long long)xs * ys >= (1ll << 31)) xs = ys = 0;
		steps_out[2 * i] = xs;
		steps_out[2 * i + 1] = ys;
	}
}

__global__ void __launch_bounds__(256) og_line_steps_kernel(
	const float2 *__restrict__ grid, int gh, int gw, int cell,
	const og_line *__restrict__ lines, int n, int32_t *__restrict__ steps_out) {
	og_line_steps_body(grid, gh, gw, cell, lines, n, steps_out, blockIdx.x * blockDim.x + threadIdx.x);
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

__device__ __forceinline__ void og_line_scan_body(
	const int32_t *steps, int n, int64_t *__restrict__ pix_offsets, int32_t *__restrict__ chunk_offsets,
	long long *ws64, int *ws32) {
	long long carry_pix = 0;
	int carry_chunks = 0;
	for (int base = 0; base < n; base += blockDim.x) {
		const int i = base + threadIdx.x;
		int xs = 0, ys = 0;
		if (i < n) {
			xs = steps[2 * i];
			ys = steps[2 * i + 1];
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
		chunk_offsets[n + 1] = 0;  // the work ticket of og_strips_screened_kernel (zero between calls)
	}
}

__global__ void __launch_bounds__(1024) og_line_scan_kernel(
	const int32_t *__restrict__ steps, int n, int64_t *__restrict__ pix_offsets, int32_t *__restrict__ chunk_offsets) {
	__shared__ long long ws64[33];
	__shared__ int ws32[33];
	og_line_scan_body(steps, n, pix_offsets, chunk_offsets, ws64, ws32);
}

// Both steps in one launch for a page's worth of lines: one cluster of eight CTAs computes the step
// counts (four lanes per line, 2048 lines per pass), a cluster barrier publishes them, and the
// first CTA runs the prefix sums.  Two dependent ~5 us launches become one.
constexpr int PLAN_CTAS = 8;
constexpr int PLAN_MAX_LINES = 8192;

__global__ void __cluster_dims__(PLAN_CTAS, 1, 1) __launch_bounds__(1024) og_line_plan_kernel(
	const float2 *__restrict__ grid, int gh, int gw, int cell,
	const og_line *__restrict__ lines, int n, int32_t *steps_out,
	int64_t *__restrict__ pix_offsets, int32_t *__restrict__ chunk_offsets) {
	__shared__ long long ws64[33];
	__shared__ int ws32[33];
	for (int base = 0; base < 4 * n; base += PLAN_CTAS * 1024)
		og_line_steps_body(grid, gh, gw, cell, lines, n, steps_out, base + blockIdx.x * 1024 + threadIdx.x);
	// release / acquire at cluster scope orders the global stores of all eight CTAs before the scan
	asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
	if (blockIdx.x == 0) og_line_scan_body(steps_out, n, pix_offsets, chunk_offsets, ws64, ws32);
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
	// MODE_STRIP (src = warped page) / MODE_LABELS (src = label map)
	const uint8_t *src;
	int H, W;
	int src_pitch;
	uint8_t *strips_out;
	int32_t *bbox_out;
	int64_t capacity;
	// MODE_LABELS
	float scale_x, scale_y;
	int n_classes;
	unsigned long long *counts_out;
	unsigned long long *overflow_out;
	// map-gather engine: derived scalars, filled in by the launcher so that the kernels read them
	// straight from the constant bank
	int Wd, Hm, Wm, W1, H1;
	unsigned one_bits;  // 0x3F800000, kept opaque to the compiler (one LOP3 instead of two per fraction)
	int *ticket;        // = chunk_offsets + n + 1
};

struct RowEntry {
	int r0, r1;  // sparse rows relative to the staged band
	float b0, b1;
};

struct ItemCtx {
	int valid;
	int li, c0, ncols, xsteps, ysteps;
	int staged, R0, C0, NR, NC, SR0, NSR;
	int rows_tabled;
	long long pix0;
	og_lin lx, ly;
	double cpx, cpy;
	double2 rowbase[MAX_ROWS];  // (ys[r] + p0), block.py:241
};

__device__ __forceinline__ int find_line(const int32_t *__restrict__ chunk_offsets, int n, int item) {
	int lo = 0, hi = n;  // invariant: chunk_offsets[lo] <= item < chunk_offsets[hi]
	while (hi - lo > 1) {
		int mid = (lo + hi) >> 1;
		if (__ldg(chunk_offsets + mid) <= item) lo = mid; else hi = mid;
	}
	return lo;
}

// executed by the whole planner warp
template <int MODE>
__device__ void plan_item(const LineParams &p, int item, ItemCtx *c) {
	const int lane = threadIdx.x & 31;
	const int Hd = p.gh * p.cell, Wd = p.gw * p.cell;
	const double inv_cell = 1.0 / (double)p.cell;
	const int li = find_line(p.chunk_offsets, p.n, item);
	const int chunk = item - __ldg(p.chunk_offsets + li);
	const int xsteps = __ldg(p.steps + 2 * li), ysteps = __ldg(p.steps + 2 * li + 1);
	const long long pix0 = (MODE != MODE_LABELS) ? __ldg(p.pix_offsets + li) : 0;
	const bool valid = MODE == MODE_LABELS || pix0 + (long long)xsteps * ysteps <= p.capacity;
	const og_line *ln = p.lines + li;
	const double cpx = ln->cp[0], cpy = ln->cp[1];
	const int c0 = chunk * CW;
	const int ncols = min(CW, xsteps - c0);
	const og_lin lx = og_linspace_setup(ln->cright[0], ln->cright[1], xsteps);
	const og_lin ly = og_linspace_setup(ln->up[0], ln->up[1], ysteps);

	// region of dense-map nodes under this chunk: the lattice is affine, so its extremes
	// are at the four corners (one margin cell absorbs rounding)
	double ymin = 1e300, ymax = -1e300, xmin = 1e300, xmax = -1e300;
#pragma unroll
	for (int k = 0; k < 4; ++k) {
		og_d2 yr = og_linspace_at(ly, (k & 1) ? ysteps - 1 : 0);
		og_d2 xc = og_linspace_at(lx, (k & 2) ? c0 + ncols - 1 : c0);
		const double px = __dadd_rn(__dadd_rn(yr.x, cpx), xc.x);
		const double py = __dadd_rn(__dadd_rn(yr.y, cpy), xc.y);
		ymin = fmin(ymin, py); ymax = fmax(ymax, py);
		xmin = fmin(xmin, px); xmax = fmax(xmax, px);
	}
	bool staged = (ymin == ymin) && (xmin == xmin) && fabs(ymin) < 1e9 && fabs(ymax) < 1e9 &&
		fabs(xmin) < 1e9 && fabs(xmax) < 1e9;
	int R0 = 0, C0 = 0, NR = 0, NC = 0, SR0 = 0, NSR = 0;
	if (staged) {
		R0 = min(max((int)floor(ymin) - 1, 0), Hd - 2);
		const int R1 = min(max((int)floor(ymax) + 1, 0), Hd - 2) + 1;
		C0 = min(max((int)floor(xmin) - 1, 0), Wd - 2);
		const int C1 = min(max((int)floor(xmax) + 1, 0), Wd - 2) + 1;
		NR = R1 - R0 + 1;
		NC = C1 - C0 + 1;
		og_axis a0 = og_resize_axis(R0, inv_cell), a1 = og_resize_axis(R1, inv_cell);
		SR0 = min(max(a0.i, 0), p.gh - 1);
		const int SR1 = min(max(a1.i + 1, 0), p.gh - 1);
		NSR = SR1 - SR0 + 1;
		const size_t need = (size_t)NR * NC * sizeof(double2) + (size_t)NSR * NC * sizeof(float2) +
			(size_t)NR * sizeof(RowEntry);
		staged = need <= (size_t)SMEM_BUDGET;
	}
	const int tabled = ysteps <= MAX_ROWS;
	if (tabled) {
		for (int r = lane; r < ysteps; r += 32) {
			og_d2 yr = og_linspace_at(ly, r);
			c->rowbase[r] = make_double2(__dadd_rn(yr.x, cpx), __dadd_rn(yr.y, cpy));
		}
	}
	if (lane == 0) {
		c->valid = valid;
		c->li = li; c->c0 = c0; c->ncols = ncols; c->xsteps = xsteps; c->ysteps = ysteps;
		c->staged = staged; c->R0 = R0; c->C0 = C0; c->NR = NR; c->NC = NC; c->SR0 = SR0; c->NSR = NSR;
		c->rows_tabled = tabled;
		c->pix0 = pix0;
		c->lx = lx; c->ly = ly;
		c->cpx = cpx; c->cpy = cpy;
	}
	__syncwarp();
}

template <int MODE, bool BBOX>
__global__ void __launch_bounds__(LTHREADS, 2) og_lines_kernel(const __grid_constant__ LineParams p) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	__shared__ ItemCtx s_ctx[2];
	__shared__ unsigned int s_hist[MODE == MODE_LABELS ? MAX_CLASSES + 1 : 1];
	__shared__ int s_bbox[4];

	const int total_items = __ldg(p.chunk_offsets + p.n);
	const int tid = threadIdx.x;
	const bool planner = tid >= WORKERS;
	const int first = blockIdx.x;
	if (first >= total_items) return;

	if (planner) plan_item<MODE>(p, first, &s_ctx[0]);
	bar_sync(0, LTHREADS);

	int k = 0;
	for (int item = first; item < total_items; item += gridDim.x, ++k) {
		if (planner) {
			const int next = item + gridDim.x;
			if (next < total_items) plan_item<MODE>(p, next, &s_ctx[(k + 1) & 1]);
			bar_sync(0, LTHREADS);
			continue;
		}
		const ItemCtx &c = s_ctx[k & 1];
		if (!c.valid) {
			bar_sync(0, LTHREADS);
			continue;
		}
		const int Hd = p.gh * p.cell, Wd = p.gw * p.cell;
		const double inv_cell = 1.0 / (double)p.cell;
		const int NR = c.NR, NC = c.NC, NSR = c.NSR, R0 = c.R0, C0 = c.C0;
		const bool staged = c.staged != 0;
		double2 *nodes = reinterpret_cast<double2 *>(smem_raw);
		float2 *hs = reinterpret_cast<float2 *>(smem_raw + (size_t)NR * NC * sizeof(double2));
		RowEntry *rows = reinterpret_cast<RowEntry *>(smem_raw + (size_t)NR * NC * sizeof(double2) +
			(size_t)NSR * NC * sizeof(float2));

		if (MODE == MODE_LABELS)
			for (int q = tid; q <= MAX_CLASSES; q += WORKERS) s_hist[q] = 0;
		if (BBOX && tid < 4) s_bbox[tid] = (tid < 2) ? INT_MAX : INT_MIN;
		if (staged) {
			// horizontal pass over the sparse rows of the band (HResizeLinear), one (row, column)
			// entry per thread
			for (int e = tid; e < NSR * NC; e += WORKERS) {
				const int r = e / NC, j = e - r * NC;
				og_axis ax = og_resize_axis_x(C0 + j, inv_cell, p.gw);
				const int c1 = min(ax.i + 1, p.gw - 1);
				const float w0 = __fsub_rn(1.f, ax.f), w1 = ax.f;
				const float2 *g = p.grid + (c.SR0 + r) * p.gw;
				const float2 s0 = __ldg(g + ax.i), s1 = __ldg(g + c1);
				hs[e] = make_float2(og_lerp_cv(s0.x, s1.x, w0, w1), og_lerp_cv(s0.y, s1.y, w0, w1));
			}
			for (int i = tid; i < NR; i += WORKERS) {
				og_axis ay = og_resize_axis(R0 + i, inv_cell);
				RowEntry e;
				e.r0 = (min(max(ay.i, 0), p.gh - 1) - c.SR0) * NC;
				e.r1 = (min(max(ay.i + 1, 0), p.gh - 1) - c.SR0) * NC;
				e.b1 = ay.f;
				e.b0 = __fsub_rn(1.f, ay.f);
				rows[i] = e;
			}
			bar_sync(1, WORKERS);
			// vertical pass (VResizeLinear) -> float64 nodes.  A thread owns one column and a
			// band of rows; the two horizontally interpolated sparse rows stay in registers
			// until the row crosses into the next sparse cell.
			{
				const int groups = max(1, WORKERS / NC);
				const int g = tid / NC, j = tid - g * NC;
				if (g < groups) {
					const int band = (NR + groups - 1) / groups;
					const int ibeg = g * band, iend = min(NR, ibeg + band);
					int cr0 = -1, cr1 = -1;
					float2 u = make_float2(0.f, 0.f), v = u;
					const int jstep = (NC <= WORKERS) ? NC * groups : WORKERS;  // NC > WORKERS: several passes
					for (int jj = j; jj < NC; jj += jstep) {
						cr0 = cr1 = -1;
						for (int i = ibeg; i < iend; ++i) {
							const RowEntry e = rows[i];
							if (e.r0 != cr0) { cr0 = e.r0; u = hs[e.r0 + jj]; }
							if (e.r1 != cr1) { cr1 = e.r1; v = hs[e.r1 + jj]; }
							nodes[i * NC + jj] = make_double2((double)og_lerp_cv(u.x, v.x, e.b0, e.b1),
								(double)og_lerp_cv(u.y, v.y, e.b0, e.b1));
						}
					}
				}
			}
		}
		bar_sync(1, WORKERS);

		// ---- sample phase: thread = (column, row group).  Software pipeline: the taps of
		// batch t are loaded before the blend of batch t-1, so their latency hides behind
		// the float64 work of the next batch.
		const int col = tid & (CW - 1), rq = tid / CW;
		const int xsteps = c.xsteps, ysteps = c.ysteps;
		float bx0 = 3.0e38f, by0 = 3.0e38f, bx1 = -3.0e38f, by1 = -3.0e38f;
		if (col < c.ncols) {
			const int cc = c.c0 + col;
			const og_d2 xc = og_linspace_at(c.lx, cc);
			const int NRm = NR - 2, NCm = NC - 2, Hm = Hd - 2, Wm = Wd - 2;
			const bool tabled = c.rows_tabled != 0;
			uint8_t *strip = (MODE == MODE_STRIP) ? p.strips_out + c.pix0 + cc : nullptr;
			float2 *gout = (MODE == MODE_GRID) ? reinterpret_cast<float2 *>(p.grids_out) + c.pix0 + cc : nullptr;
			const uint8_t *__restrict__ src = p.src;
			const int H = p.H, W = p.W, spitch = p.src_pitch;
			og_taps pend[RB];
			int pend_row[RB];
			bool have_pend = false;
			for (int rbase = rq; rbase < ysteps; rbase += RB * RSPLIT) {
				float gx[RB], gy[RB];
				int row[RB];
#pragma unroll
				for (int u = 0; u < RB; ++u) {
					const int r = rbase + u * RSPLIT;
					row[u] = r < ysteps ? r : -1;
					const int rr = r < ysteps ? r : rbase;
					double2 rb;
					if (tabled) {
						rb = c.rowbase[rr];
					} else {
						og_d2 yr = og_linspace_at(c.ly, rr);
						rb = make_double2(__dadd_rn(yr.x, c.cpx), __dadd_rn(yr.y, c.cpy));
					}
					// dewarped_grid = (ys + p0)[:, None] + xs[None, :]   (block.py:241)
					const double px = __dadd_rn(rb.x, xc.x);
					const double py = __dadd_rn(rb.y, xc.y);
					og_d2 w;
					if (staged) {
						// i = clip(floor(y), 0, n - 2) (SciPy find_indices); |y| < 1e9 here
						const int i0 = min(max(__double2int_rd(py), 0), Hm);
						const int i1 = min(max(__double2int_rd(px), 0), Wm);
						const int ri = i0 - R0, ci = i1 - C0;
						// the staged region covers the chunk by construction; the clamp keeps the
						// shared-memory reads in bounds and `ok` sends a (never observed) outlier
						// through the unstaged path below
						const bool ok = (unsigned)ri <= (unsigned)NRm && (unsigned)ci <= (unsigned)NCm;
						const int rs = min(max(ri, 0), NRm), cs = min(max(ci, 0), NCm);
						const double a = __dsub_rn(py, (double)i0), b = __dsub_rn(px, (double)i1);
						const double wa = __dsub_rn(1.0, a), wb = __dsub_rn(1.0, b);
						const double w00 = __dmul_rn(wa, wb), w01 = __dmul_rn(wa, b);
						const double w10 = __dmul_rn(a, wb), w11 = __dmul_rn(a, b);
						const double2 *nd = nodes + (rs * NC + cs);
						const double2 v00 = nd[0], v01 = nd[1], v10 = nd[NC], v11 = nd[NC + 1];
						w.x = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(v00.x, w00), __dmul_rn(v01.x, w01)),
							__dmul_rn(v10.x, w10)), __dmul_rn(v11.x, w11));
						w.y = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(v00.y, w00), __dmul_rn(v01.y, w01)),
							__dmul_rn(v10.y, w10)), __dmul_rn(v11.y, w11));
						if (!ok) w = og_inverse_point(p.grid, p.gh, p.gw, p.cell, inv_cell, py, px);
					} else {
						w = og_inverse_point(p.grid, p.gh, p.gw, p.cell, inv_cell, py, px);
					}
					gx[u] = __double2float_rn(w.x);
					gy[u] = __double2float_rn(w.y);
				}
				if (MODE == MODE_GRID) {
#pragma unroll
					for (int u = 0; u < RB; ++u)
						if (row[u] >= 0) gout[row[u] * xsteps] = make_float2(gx[u], gy[u]);
				} else if (MODE == MODE_STRIP) {
					og_taps cur[RB];
					og_load_taps_batch<RB>(src, H, W, spitch, gx, gy, cur);
					if (have_pend) {
#pragma unroll
						for (int u = 0; u < RB; ++u)
							if (pend_row[u] >= 0)  // pixels[::-1, :]  (block.py:257)
								strip[(ysteps - 1 - pend_row[u]) * xsteps] = (uint8_t)og_blend_taps(pend[u]);
					}
#pragma unroll
					for (int u = 0; u < RB; ++u) {
						pend[u] = cur[u];
						pend_row[u] = row[u];
						if (BBOX) {  // a dead slot repeats a live row's coordinates: harmless for min/max
							bx0 = fminf(bx0, gx[u]); bx1 = fmaxf(bx1, gx[u]);
							by0 = fminf(by0, gy[u]); by1 = fmaxf(by1, gy[u]);
						}
					}
					have_pend = true;
				} else {
#pragma unroll
					for (int u = 0; u < RB; ++u) {
						if (row[u] >= 0) {
							// scale_grid (batch/detect/lines.py:18-22) then cv2.remap INTER_NEAREST
							const float sx = __fmul_rn(gx[u], p.scale_x), sy = __fmul_rn(gy[u], p.scale_y);
							const int lab = og_sample_nearest_fast(src, H, W, spitch, sx, sy);
							atomicAdd(&s_hist[min(lab, p.n_classes)], 1u);
						}
					}
				}
			}
			if (MODE == MODE_STRIP && have_pend) {
#pragma unroll
				for (int u = 0; u < RB; ++u)
					if (pend_row[u] >= 0)
						strip[(ysteps - 1 - pend_row[u]) * xsteps] = (uint8_t)og_blend_taps(pend[u]);
			}
		}
		if (BBOX) {
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
			bar_sync(1, WORKERS);
			if (tid < 4 && s_bbox[0] <= s_bbox[2]) {
				if (tid < 2) atomicMin(p.bbox_out + 4 * c.li + tid, s_bbox[tid]);
				else atomicMax(p.bbox_out + 4 * c.li + tid, s_bbox[tid]);
			}
		}
		if (MODE == MODE_LABELS) {
			bar_sync(1, WORKERS);
			for (int q = tid; q < p.n_classes; q += WORKERS)
				if (s_hist[q]) atomicAdd(p.counts_out + (size_t)c.li * p.n_classes + q, (unsigned long long)s_hist[q]);
			if (tid == 0 && s_hist[p.n_classes] && p.overflow_out != nullptr)
				atomicAdd(p.overflow_out + c.li, (unsigned long long)s_hist[p.n_classes]);
		}
		bar_sync(0, LTHREADS);
	}
}

// ---------------------------------------------------------------------------------------
// Map-gather engine: when the caller holds the materialised dense map (the page kernel
// emits it, as the reference's lru_cached Grid.points("full") does), the four nodes of a
// sample are read straight from it.  No staging, no shared-memory nodes, no CTA barriers:
// a warp owns a (line, 32-column) item, a lane walks one strip column; node loads and tap
// loads of two consecutive rows are in flight together and the byte blend of the previous
// pair is deferred behind them.
// ---------------------------------------------------------------------------------------
constexpr int MW = 32;            // strip columns per warp item
constexpr int MWARPS = 8;         // warps per CTA
constexpr int MTHREADS = MWARPS * 32;
constexpr int MAP_OFFS_CAP = 4096;  // line prefix sums cached in shared memory up to this many lines
#ifndef OG_MAP_MINBLOCKS
#define OG_MAP_MINBLOCKS 3
#endif

#ifndef OG_LINES_SCREEN
#define OG_LINES_SCREEN 1
#endif

// One sample of the dense map in the reference's float64 order (SciPy _evaluate_linear), cast to
// float32 and quantised as cv2.remap does: the slow, always-valid path behind the screening below.
__device__ __forceinline__ float2 og_map_sample_exact(const float2 *__restrict__ map, int Wd, int Hm, int Wm, double px, double py) {
	const int i0 = min(max(__double2int_rd(py), 0), Hm);
	const int i1 = min(max(__double2int_rd(px), 0), Wm);
	const double a = __dsub_rn(py, (double)i0), b = __dsub_rn(px, (double)i1);
	const float2 *m = map + (i0 * Wd + i1);
	const float2 v00 = __ldg(m), v01 = __ldg(m + 1), v10 = __ldg(m + Wd), v11 = __ldg(m + Wd + 1);
	const double wa = __dsub_rn(1.0, a), wb = __dsub_rn(1.0, b);
	const double w00 = __dmul_rn(wa, wb), w01 = __dmul_rn(wa, b);
	const double w10 = __dmul_rn(a, wb), w11 = __dmul_rn(a, b);
	const double wx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00.x, w00), __dmul_rn((double)v01.x, w01)),
		__dmul_rn((double)v10.x, w10)), __dmul_rn((double)v11.x, w11));
	const double wy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00.y, w00), __dmul_rn((double)v01.y, w01)),
		__dmul_rn((double)v10.y, w10)), __dmul_rn((double)v11.y, w11));
	return make_float2(__double2float_rn(wx), __double2float_rn(wy));
}

__device__ __noinline__ int2 og_map_sample_exact_fixed(const float2 *__restrict__ map, int Wd, int Hm, int Wm, double px, double py) {
	const float2 g = og_map_sample_exact(map, Wd, Hm, Wm, px, py);
	return make_int2(__float2int_rn(__fmul_rn(g.x, 32.f)), __float2int_rn(__fmul_rn(g.y, 32.f)));
}

// Screening constants.  OG_FIX_MAGIC = 1.5 * 2^29: a double in [2^29, 2^30) has an ulp of 2^-23,
// so (v + OG_FIX_MAGIC) rounded DOWN holds floor(v * 2^23) in its mantissa: the integer part
// of v in bits 23.. and 23 fraction bits below, without a conversion instruction.
// OG_RINT_MAGIC = 1.5 * 2^23: fl32(32 x + OG_RINT_MAGIC) holds cvRound(32 x) in its low mantissa bits.
#define OG_FIX_MAGIC 805306368.0
#define OG_RINT_MAGIC 12582912.0f
#define OG_RINT_BITS 0x4B400000

// ---------------------------------------------------------------------------------------
// Screened strip sampling.  What cv2.remap uses of a sample is cvRound(32 * fl32(X)), with X
// SciPy's float64 blend of the four dense-map nodes.  The blend is evaluated in float32 as
// v00 + delta with |delta - exact| <= 26 u G (u = 2^-24, G = the largest node difference of the
// cell; derivation in DESIGN.md), so with eps = 32 u G + 2^-30 the exact X lies in
// [v00 + delta - eps, v00 + delta + eps].  Rounding to float32 and cvRound are monotone: when
// both ends of the interval quantise to the same fixed-point value, that IS the reference's
// value.  Otherwise (a few samples in 10^4), and for samples outside the interior of the map or
// of the page, the float64 path decides, so the strips stay bit-identical.
// ---------------------------------------------------------------------------------------
struct og_nodes {
	float2 v00, v01, v10, v11;
	float a, b;
	bool ok;
};

// stage 1: cell index + fractions of (px, py) without a conversion instruction, node loads.
// The caller guarantees |px|, |py| < 2^27, so px + OG_FIX_MAGIC stays in [2^29, 2^30).
__device__ __forceinline__ void og_nodes_load(og_nodes &n, double px, double py,
	const float2 *__restrict__ map, int Wd, int Hm, int Wm, unsigned one_bits) {
	const double tx = __dadd_rd(px, OG_FIX_MAGIC), ty = __dadd_rd(py, OG_FIX_MAGIC);
	const int txl = __double2loint(tx), tyl = __double2loint(ty);
	// mantissa = 2^51 + floor(v * 2^23); the three bits above it belong to the exponent 0x41C
	const unsigned i1 = __funnelshift_r(txl, __double2hiint(tx), 23) - 0x90000000u;
	const unsigned i0 = __funnelshift_r(tyl, __double2hiint(ty), 23) - 0x90000000u;
	n.ok = i0 <= (unsigned)Hm && i1 <= (unsigned)Wm;  // a negative floor wraps to a huge index
	n.a = __fsub_rn(__int_as_float((tyl & 0x7FFFFF) | one_bits), 1.f);
	n.b = __fsub_rn(__int_as_float((txl & 0x7FFFFF) | one_bits), 1.f);
	const float2 *m = map + (n.ok ? i0 * (unsigned)Wd + i1 : 0u);
	n.v00 = __ldg(m);
	n.v10 = __ldg(m + Wd);
#ifdef OG_EXP_NODES2  // experiment (wrong pixels): how much of the time is the number of node loads?
	n.v01 = n.v00;
	n.v11 = n.v10;
#else
	n.v01 = __ldg(m + 1);
	n.v11 = __ldg(m + Wd + 1);
#endif
}

// stage 2: float32 interval of the blend -> fixed-point coordinates; false when undecided
__device__ __forceinline__ bool og_nodes_screen(const og_nodes &n, int W1, int H1, int &sx, int &sy) {
	const float d01x = __fsub_rn(n.v01.x, n.v00.x), d10x = __fsub_rn(n.v10.x, n.v00.x), d11x = __fsub_rn(n.v11.x, n.v10.x);
	const float d01y = __fsub_rn(n.v01.y, n.v00.y), d10y = __fsub_rn(n.v10.y, n.v00.y), d11y = __fsub_rn(n.v11.y, n.v10.y);
	const float g = fmaxf(fmaxf(fmaxf(fabsf(d01x), fabsf(d10x)), fmaxf(fabsf(d11x), fabsf(d01y))), fmaxf(fabsf(d10y), fabsf(d11y)));
	const float eps = __fmaf_rn(g, 0x1p-19f, 0x1p-30f);
	const float dx = __fmaf_rn(n.a, __fmaf_rn(n.b, __fsub_rn(d11x, d01x), d10x), __fmul_rn(n.b, d01x));
	const float dy = __fmaf_rn(n.a, __fmaf_rn(n.b, __fsub_rn(d11y, d01y), d10y), __fmul_rn(n.b, d01y));
	const float uxl = __fmaf_rn(__fadd_rn(n.v00.x, __fsub_rn(dx, eps)), 32.f, OG_RINT_MAGIC);
	const float uxh = __fmaf_rn(__fadd_rn(n.v00.x, __fadd_rn(dx, eps)), 32.f, OG_RINT_MAGIC);
	const float uyl = __fmaf_rn(__fadd_rn(n.v00.y, __fsub_rn(dy, eps)), 32.f, OG_RINT_MAGIC);
	const float uyh = __fmaf_rn(__fadd_rn(n.v00.y, __fadd_rn(dy, eps)), 32.f, OG_RINT_MAGIC);
	sx = __float_as_int(uxl) - OG_RINT_BITS;
	sy = __float_as_int(uyl) - OG_RINT_BITS;
	// the mantissa only holds cvRound(32 x) while 32 x stays below 2^22: the interior test keeps
	// x inside the page (< 32767), and a NaN compares unequal
	return n.ok && uxl == uxh && uyl == uyh && (unsigned)(sx >> 5) < (unsigned)W1 && (unsigned)(sy >> 5) < (unsigned)H1;
}

// the undecided sample, start to finish in the reference's arithmetic
__device__ __noinline__ og_taps og_sample_taps_exact(const uint8_t *__restrict__ src, int H, int W, int pitch,
	const float2 *__restrict__ map, int Wd, int Hm, int Wm, double px, double py) {
	const int2 s = og_map_sample_exact_fixed(map, Wd, Hm, Wm, px, py);
	const int sx[1] = {s.x}, sy[1] = {s.y};
	og_taps t[1];
	og_load_taps_fixed<1>(src, H, W, pitch, sx, sy, t);
	return t[0];
}

__device__ __forceinline__ void og_taps_interior(og_taps &t, const uint8_t *__restrict__ src, int pitch, int sx, int sy) {
	const uint8_t *q = src + ((sy >> 5) * pitch + (sx >> 5));
	t.fx = sx & 31;
	t.fy = sy & 31;
	t.p00 = __ldg(q);
	t.p10 = __ldg(q + pitch);
#ifdef OG_EXP_TAPS2  // experiment (wrong pixels): how much of the time is the number of tap loads?
	t.p01 = t.p00;
	t.p11 = t.p10;
#else
	t.p01 = __ldg(q + 1);
	t.p11 = __ldg(q + pitch + 1);
#endif
}

// --- label sampling (A15) through the same screening: what the NEAREST lookup uses of a sample is
// cvRound(fl32(fl32(X) * scale)) per coordinate, again monotone in X.
__device__ __forceinline__ bool og_nodes_screen_nearest(const og_nodes &n, float scx, float scy, int &ix, int &iy) {
	const float d01x = __fsub_rn(n.v01.x, n.v00.x), d10x = __fsub_rn(n.v10.x, n.v00.x), d11x = __fsub_rn(n.v11.x, n.v10.x);
	const float d01y = __fsub_rn(n.v01.y, n.v00.y), d10y = __fsub_rn(n.v10.y, n.v00.y), d11y = __fsub_rn(n.v11.y, n.v10.y);
	const float g = fmaxf(fmaxf(fmaxf(fabsf(d01x), fabsf(d10x)), fmaxf(fabsf(d11x), fabsf(d01y))), fmaxf(fabsf(d10y), fabsf(d11y)));
	const float eps = __fmaf_rn(g, 0x1p-19f, 0x1p-30f);
	const float dx = __fmaf_rn(n.a, __fmaf_rn(n.b, __fsub_rn(d11x, d01x), d10x), __fmul_rn(n.b, d01x));
	const float dy = __fmaf_rn(n.a, __fmaf_rn(n.b, __fsub_rn(d11y, d01y), d10y), __fmul_rn(n.b, d01y));
	// scale_grid (batch/detect/lines.py:18-22) rounds the product to float32 before cv2.remap rounds
	// it to an integer: a multiply, then the magic add
	const float uxl = __fadd_rn(__fmul_rn(__fadd_rn(n.v00.x, __fsub_rn(dx, eps)), scx), OG_RINT_MAGIC);
	const float uxh = __fadd_rn(__fmul_rn(__fadd_rn(n.v00.x, __fadd_rn(dx, eps)), scx), OG_RINT_MAGIC);
	const float uyl = __fadd_rn(__fmul_rn(__fadd_rn(n.v00.y, __fsub_rn(dy, eps)), scy), OG_RINT_MAGIC);
	const float uyh = __fadd_rn(__fmul_rn(__fadd_rn(n.v00.y, __fadd_rn(dy, eps)), scy), OG_RINT_MAGIC);
	ix = __float_as_int(uxl) - OG_RINT_BITS;
	iy = __float_as_int(uyl) - OG_RINT_BITS;
	// beyond |x| = 2^22 the mantissa no longer holds the integer, but then (ix, iy) cannot land
	// inside the label map either; a NaN compares unequal
	return n.ok && uxl == uxh && uyl == uyh;
}

__device__ __noinline__ int og_sample_label_exact(const uint8_t *__restrict__ src, int H, int W, int pitch,
	const float2 *__restrict__ map, int Wd, int Hm, int Wm, double px, double py, float scx, float scy) {
	const float2 g = og_map_sample_exact(map, Wd, Hm, Wm, px, py);
	return og_sample_nearest_fast(src, H, W, pitch, __fmul_rn(g.x, scx), __fmul_rn(g.y, scy));
}

// One strip column: `rows` rows whose (ys + p0) sit in the warp's shared-memory table at
// tab_saddr; `out` points at the column's pixel of the FIRST row and moves up the flipped strip
// (pixels[::-1, :], block.py:257).  Three batches of RB rows are in flight: node loads of batch
// k+1, tap loads of batch k, blend + store of batch k-1.
template <int NB>
__device__ __forceinline__ void og_strip_column_screened(const LineParams &p, const float2 *__restrict__ map,
	unsigned tab_saddr, int rows, og_d2 xc, uint8_t *__restrict__ out, int xsteps) {
	const uint8_t *__restrict__ src = p.src;
	auto point = [&](int r, double &px, double &py) {
		double rbx, rby;
		asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rbx), "=d"(rby) : "r"(tab_saddr + r * 16));
		px = __dadd_rn(rbx, xc.x);  // dewarped_grid = (ys + p0)[:, None] + xs[None, :]   (block.py:241)
		py = __dadd_rn(rby, xc.y);
	};
	auto load = [&](og_nodes (&n)[NB], int r) {
#pragma unroll
		for (int u = 0; u < NB; ++u) {
			double px, py;
			point(r + u, px, py);
			og_nodes_load(n[u], px, py, map, p.Wd, p.Hm, p.Wm, p.one_bits);
		}
	};
	auto taps = [&](const og_nodes (&n)[NB], og_taps (&t)[NB], int r) {
		int sx[NB], sy[NB];
		bool ok[NB], good = true;
#pragma unroll
		for (int u = 0; u < NB; ++u) {
			ok[u] = og_nodes_screen(n[u], p.W1, p.H1, sx[u], sy[u]);
			good = good && ok[u];
		}
		if (good) {
#pragma unroll
			for (int u = 0; u < NB; ++u) og_taps_interior(t[u], src, p.src_pitch, sx[u], sy[u]);
		} else {
#pragma unroll
			for (int u = 0; u < NB; ++u) {
				if (ok[u]) {
					og_taps_interior(t[u], src, p.src_pitch, sx[u], sy[u]);
				} else {
					double px, py;
					point(r + u, px, py);
					t[u] = og_sample_taps_exact(src, p.H, p.W, p.src_pitch, map, p.Wd, p.Hm, p.Wm, px, py);
				}
			}
		}
	};
	auto store = [&](const og_taps (&t)[NB]) {
#pragma unroll
		for (int u = 0; u < NB; ++u) {
			*out = (uint8_t)og_blend_taps(t[u]);
			out -= xsteps;
		}
	};

	const int nb = rows / NB;  // full batches; the caller handles rows % NB with NB = 1
	if (nb == 0) return;
	og_nodes n0[NB], n1[NB];
	og_taps t0[NB], t1[NB];
	load(n0, 0);
	load(n1, min(1, nb - 1) * NB);
	taps(n0, t0, 0);
	int b = 1;
	for (; b + 1 < nb; b += 2) {  // on entry: t0 = taps of batch b-1, n1 = nodes of batch b
		load(n0, (b + 1) * NB);
		store(t0);  // before the next taps are requested: only one tap buffer is live at a time
		taps(n1, t1, b * NB);
		load(n1, min(b + 2, nb - 1) * NB);
		store(t1);
		taps(n0, t0, (b + 1) * NB);
	}
	store(t0);
	if (b < nb) {
		taps(n1, t1, b * NB);
		store(t1);
	}
}

// the same column in the reference's arithmetic only (lines with absurd coordinates)
__device__ __noinline__ void og_strip_column_exact(const LineParams &p, const float2 *__restrict__ map,
	unsigned tab_saddr, int rows, og_d2 xc, uint8_t *__restrict__ out, int xsteps) {
	for (int r = 0; r < rows; ++r) {
		double rbx, rby;
		asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rbx), "=d"(rby) : "r"(tab_saddr + r * 16));
		const og_taps t = og_sample_taps_exact(p.src, p.H, p.W, p.src_pitch, map, p.Wd, p.Hm, p.Wm,
			__dadd_rn(rbx, xc.x), __dadd_rn(rby, xc.y));
		*out = (uint8_t)og_blend_taps(t);
		out -= xsteps;
	}
}

// One column of a line's label samples (A15): as og_strip_column_screened, with one NEAREST byte
// per sample instead of four taps and a shared-memory histogram instead of the strip.
template <int NB>
__device__ __forceinline__ void og_label_column_screened(const LineParams &p, const float2 *__restrict__ map,
	unsigned tab_saddr, int rows, og_d2 xc, unsigned int *hist) {
	const uint8_t *__restrict__ src = p.src;
	auto point = [&](int r, double &px, double &py) {
		double rbx, rby;
		asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rbx), "=d"(rby) : "r"(tab_saddr + r * 16));
		px = __dadd_rn(rbx, xc.x);
		py = __dadd_rn(rby, xc.y);
	};
	auto load = [&](og_nodes (&n)[NB], int r) {
#pragma unroll
		for (int u = 0; u < NB; ++u) {
			double px, py;
			point(r + u, px, py);
			og_nodes_load(n[u], px, py, map, p.Wd, p.Hm, p.Wm, p.one_bits);
		}
	};
	auto fetch = [&](const og_nodes (&n)[NB], int (&lab)[NB], int r) {
#pragma unroll
		for (int u = 0; u < NB; ++u) {
			int ix, iy;
			if (og_nodes_screen_nearest(n[u], p.scale_x, p.scale_y, ix, iy)) {
				// cv2.remap INTER_NEAREST: outside the label map -> BORDER_CONSTANT 0
				lab[u] = ((unsigned)ix < (unsigned)p.W && (unsigned)iy < (unsigned)p.H) ? (int)__ldg(src + (iy * p.src_pitch + ix)) : 0;
			} else {
				double px, py;
				point(r + u, px, py);
				lab[u] = og_sample_label_exact(src, p.H, p.W, p.src_pitch, map, p.Wd, p.Hm, p.Wm, px, py, p.scale_x, p.scale_y);
			}
		}
	};
	auto count = [&](const int (&lab)[NB]) {
#pragma unroll
		for (int u = 0; u < NB; ++u) atomicAdd(hist + min(lab[u], p.n_classes), 1u);
	};

	const int nb = rows / NB;
	if (nb == 0) return;
	og_nodes n0[NB], n1[NB];
	int l0[NB], l1[NB];
	load(n0, 0);
	load(n1, min(1, nb - 1) * NB);
	fetch(n0, l0, 0);
	int b = 1;
	for (; b + 1 < nb; b += 2) {  // on entry: l0 = labels of batch b-1, n1 = nodes of batch b
		load(n0, (b + 1) * NB);
		count(l0);
		fetch(n1, l1, b * NB);
		load(n1, min(b + 2, nb - 1) * NB);
		count(l1);
		fetch(n0, l0, (b + 1) * NB);
	}
	count(l0);
	if (b < nb) {
		fetch(n1, l1, b * NB);
		count(l1);
	}
}

__device__ __noinline__ void og_label_column_exact(const LineParams &p, const float2 *__restrict__ map,
	unsigned tab_saddr, int rows, og_d2 xc, unsigned int *hist) {
	for (int r = 0; r < rows; ++r) {
		double rbx, rby;
		asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rbx), "=d"(rby) : "r"(tab_saddr + r * 16));
		const int lab = og_sample_label_exact(p.src, p.H, p.W, p.src_pitch, map, p.Wd, p.Hm, p.Wm,
			__dadd_rn(rbx, xc.x), __dadd_rn(rby, xc.y), p.scale_x, p.scale_y);
		atomicAdd(hist + min(lab, p.n_classes), 1u);
	}
}

// The line of a work item: 32-ary search of the prefix sums by the whole warp (three probes for
// 2000 lines instead of eleven dependent loads).  offs[0] = 0 <= c < offs[n].
__device__ __forceinline__ int og_find_line(const int32_t *offs, int n, int c, int lane) {
	int lo = 0, hi = n;  // invariant: offs[lo] <= c < offs[hi]
	while (hi - lo > 1) {
		const int step = (hi - lo + 31) >> 5;
		const int idx = lo + lane * step;
		const bool le = idx < hi && offs[idx] <= c;  // true for a prefix of the lanes, lane 0 included
		const int cnt = __popc(__ballot_sync(0xffffffffu, le));
		lo += (cnt - 1) * step;
		hi = min(lo + step, hi);
	}
	return lo;
}

// og_linspace_setup for the x and the y direction of a line at once: the four float64 divisions
// run in four lanes side by side instead of one after the other in every lane.
__device__ __forceinline__ void og_linspace_setup_pair(double crx, double cry, double upx, double upy,
	int xsteps, int ysteps, int lane, og_lin &lx, og_lin &ly) {
	const int k = lane & 3;
	const double stop = k == 0 ? crx : k == 1 ? cry : k == 2 ? upx : upy;
	const int num = k < 2 ? xsteps : ysteps;
	const double q = num > 1 ? __ddiv_rn(stop, (double)(num - 1)) : stop;
	lx.num = xsteps; lx.stopx = crx; lx.stopy = cry; lx.div = (double)(xsteps - 1);
	lx.stepx = __shfl_sync(0xffffffffu, q, 0);
	lx.stepy = __shfl_sync(0xffffffffu, q, 1);
	lx.any_zero = xsteps > 1 && (lx.stepx == 0.0 || lx.stepy == 0.0);
	ly.num = ysteps; ly.stopx = upx; ly.stopy = upy; ly.div = (double)(ysteps - 1);
	ly.stepx = __shfl_sync(0xffffffffu, q, 2);
	ly.stepy = __shfl_sync(0xffffffffu, q, 3);
	ly.any_zero = ysteps > 1 && (ly.stepx == 0.0 || ly.stepy == 0.0);
}

// Strips through the screened path: a warp owns a (line, 32-column) item, a lane one column.
// Items are dealt round-robin, so the warps that run together work on neighbouring items and
// share the DRAM pages / L2 lines of the map rows under them (one contiguous range of items per
// warp was measured: twice the DRAM traffic, 314 vs 290 us).
#ifndef OG_SCR_WARPS
#define OG_SCR_WARPS 8
#endif
#ifndef OG_SCR_TAIL_HALVES
#define OG_SCR_TAIL_HALVES 2  // items cut into row quarters at the end of the ticket list, in half rounds
#endif
#ifndef OG_SCR_BLOCKS
#define OG_SCR_BLOCKS 4
#endif
constexpr int SWARPS = OG_SCR_WARPS;  // warps per CTA: there is no CTA-wide barrier in the item loop
constexpr int STHREADS = SWARPS * 32;

template <int MODE>  // MODE_STRIP or MODE_LABELS
__global__ void __launch_bounds__(STHREADS, OG_SCR_BLOCKS) og_strips_screened_kernel(const __grid_constant__ LineParams p, const float2 *__restrict__ map) {
	__shared__ double2 s_rowbase[SWARPS][MAX_ROWS];
	__shared__ int s_chunk_offsets[MAP_OFFS_CAP];
	__shared__ unsigned int s_whist[MODE == MODE_LABELS ? SWARPS : 1][MODE == MODE_LABELS ? MAX_CLASSES + 1 : 1];
	const bool offs_cached = p.n + 1 <= MAP_OFFS_CAP;
	if (offs_cached) {
		for (int i = threadIdx.x; i <= p.n; i += STHREADS) s_chunk_offsets[i] = __ldg(p.chunk_offsets + i);
		__syncthreads();
	}
	const int32_t *chunk_offsets = offs_cached ? s_chunk_offsets : p.chunk_offsets;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int total_items = 2 * chunk_offsets[p.n];  // two 32-column halves per 64-column chunk
	const int nwarps = gridDim.x * SWARPS;
	double2 *rowbase = s_rowbase[wid];
	const unsigned rb_saddr = (unsigned)__cvta_generic_to_shared(rowbase);

	// Work is handed out through a ticket counter (the extra entry behind the prefix sums): the SMs
	// differ by +-13 % in speed on this kernel, which a static deal turns into idle time.  Tickets
	// are items in order, so the warps that run together still work on neighbouring items; the last
	// items (two quarters per warp) are cut into row quarters so that everybody finishes together.  A
	// warp draws its next ticket before it starts on the current one (the atomic's latency hides
	// behind the item), and the warp that draws the very last ticket resets the counter.
	constexpr int PIECES = 4;
	const int tail_items = min(total_items, nwarps * OG_SCR_TAIL_HALVES / 2), head_items = total_items - tail_items;
	const int total_tickets = head_items + tail_items * PIECES;
	auto draw = [&]() {
		int t = 0;
		if (lane == 0) t = atomicAdd(p.ticket, 1);
		return __shfl_sync(0xffffffffu, t, 0);
	};
	auto work = [&](int tk) {
		int item = tk, piece = 0, npieces = 1;
		if (tk >= head_items) {
			item = head_items + (tk - head_items) / PIECES;
			piece = (tk - head_items) % PIECES;
			npieces = PIECES;
		}
		const int chunk_item = item >> 1;
		const int li = og_find_line(chunk_offsets, p.n, chunk_item, lane);
		const int xsteps = __ldg(p.steps + 2 * li), ysteps = __ldg(p.steps + 2 * li + 1);
		const int c0 = (chunk_item - chunk_offsets[li]) * CW + (item & 1) * MW;
		if (c0 >= xsteps) return;
		const long long pix0 = MODE == MODE_STRIP ? __ldg(p.pix_offsets + li) : 0;
		if (MODE == MODE_STRIP && pix0 + (long long)xsteps * ysteps > p.capacity) return;
		// this warp's rows of the item (an even count, so that only the strip's own last row can be odd)
		const int row_a = (int)((long long)ysteps * piece / npieces) & ~1;
		const int row_z = piece + 1 == npieces ? ysteps : (int)((long long)ysteps * (piece + 1) / npieces) & ~1;
		if (row_a >= row_z) return;
		const og_line *ln = p.lines + li;
		const double cpx = ln->cp[0], cpy = ln->cp[1];
		const double crx = ln->cright[0], cry = ln->cright[1], upx = ln->up[0], upy = ln->up[1];
		og_lin lx, ly;
		og_linspace_setup_pair(crx, cry, upx, upy, xsteps, ysteps, lane, lx, ly);
		// every sample is cp + s * cright + t * up with s, t in [0, 1]: below this bound the
		// fixed-point trick of og_nodes_load is valid for the whole line (false for NaN / inf)
		const bool sane = fabs(cpx) + fabs(cpy) + fabs(crx) + fabs(cry) + fabs(upx) + fabs(upy) < 134217728.0;
		const int cc = c0 + lane;
		const og_d2 xc = og_linspace_at(lx, min(cc, xsteps - 1));
		if (MODE == MODE_LABELS) {
			for (int q = lane; q <= MAX_CLASSES; q += 32) s_whist[wid][q] = 0;
		}
		for (int rb0 = row_a; rb0 < row_z; rb0 += MAX_ROWS) {
			const int rows = min(row_z - rb0, MAX_ROWS);
			__syncwarp();
			for (int r = lane; r < rows; r += 32) {  // (ys + p0) of the block's rows   (block.py:241)
				const og_d2 yr = og_linspace_at(ly, rb0 + r);
				rowbase[r] = make_double2(__dadd_rn(yr.x, cpx), __dadd_rn(yr.y, cpy));
			}
			__syncwarp();
			if (cc >= xsteps) continue;
			if (MODE == MODE_LABELS) {
				unsigned int *hist = s_whist[wid];
				if (sane) {
					og_label_column_screened<RB>(p, map, rb_saddr, rows, xc, hist);
					const int done = rows / RB * RB;
					if (done < rows) og_label_column_screened<1>(p, map, rb_saddr + done * 16, rows - done, xc, hist);
				} else {
					og_label_column_exact(p, map, rb_saddr, rows, xc, hist);
				}
				continue;
			}
			uint8_t *out = p.strips_out + pix0 + cc + (long long)(ysteps - 1 - rb0) * xsteps;
			if (sane) {
				og_strip_column_screened<RB>(p, map, rb_saddr, rows, xc, out, xsteps);
				const int done = rows / RB * RB;
				if (done < rows)
					og_strip_column_screened<1>(p, map, rb_saddr + done * 16, rows - done, xc, out - (long long)done * xsteps, xsteps);
			} else {
				og_strip_column_exact(p, map, rb_saddr, rows, xc, out, xsteps);
			}
		}
		if (MODE == MODE_LABELS) {  // this ticket's share of the line's counts
			__syncwarp();
			for (int q = lane; q < p.n_classes; q += 32)
				if (s_whist[wid][q]) atomicAdd(p.counts_out + (size_t)li * p.n_classes + q, (unsigned long long)s_whist[wid][q]);
			if (lane == 0 && s_whist[wid][p.n_classes] && p.overflow_out != nullptr)
				atomicAdd(p.overflow_out + li, (unsigned long long)s_whist[wid][p.n_classes]);
			__syncwarp();
		}
	};

	int tk = draw();
	while (tk < total_tickets) {
		const int tk_next = draw();
		work(tk);
		tk = tk_next;
	}
	// every warp draws exactly one ticket past the end
	if (lane == 0 && tk == total_tickets + nwarps - 1) *p.ticket = 0;
}

template <int MODE, bool BBOX>
__global__ void __launch_bounds__(MTHREADS, OG_MAP_MINBLOCKS) og_lines_map_kernel(const __grid_constant__ LineParams p, const float2 *__restrict__ map) {
	__shared__ double2 s_rowbase[MWARPS][MAX_ROWS];
	__shared__ unsigned int s_whist[MODE == MODE_LABELS ? MWARPS : 1][MODE == MODE_LABELS ? MAX_CLASSES + 1 : 1];

	// the work-item -> line lookup is a binary search per item: keep the prefix sums in shared
	// memory when they fit, so its 11 dependent loads cost 30 instead of 300+ cycles each
	__shared__ int s_chunk_offsets[MAP_OFFS_CAP];
	const bool offs_cached = p.n + 1 <= MAP_OFFS_CAP;
	if (offs_cached) {
		for (int i = threadIdx.x; i <= p.n; i += MTHREADS) s_chunk_offsets[i] = __ldg(p.chunk_offsets + i);
		__syncthreads();
	}
	const int32_t *chunk_offsets = offs_cached ? s_chunk_offsets : p.chunk_offsets;
	const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	const int total_items = 2 * chunk_offsets[p.n];  // two 32-column halves per 64-column chunk
	const int nwarps = gridDim.x * MWARPS;
	const int Hd = p.gh * p.cell, Wd = p.gw * p.cell;
	const int Hm = Hd - 2, Wm = Wd - 2;
	const uint8_t *__restrict__ src = p.src;
	const int H = p.H, W = p.W, spitch = p.src_pitch;
	double2 *rowbase = s_rowbase[wid];
	const unsigned rb_saddr = (unsigned)__cvta_generic_to_shared(rowbase);

	for (int item = blockIdx.x * MWARPS + wid; item < total_items; item += nwarps) {
		const int chunk_item = item >> 1;
		int li = 0;
		{
			int hi = p.n;  // invariant: chunk_offsets[li] <= chunk_item < chunk_offsets[hi]
			while (hi - li > 1) {
				const int mid = (li + hi) >> 1;
				if (chunk_offsets[mid] <= chunk_item) li = mid; else hi = mid;
			}
		}
		const int xsteps = __ldg(p.steps + 2 * li), ysteps = __ldg(p.steps + 2 * li + 1);
		const int c0 = (chunk_item - chunk_offsets[li]) * CW + (item & 1) * MW;
		if (c0 >= xsteps) continue;
		const long long pix0 = (MODE != MODE_LABELS) ? __ldg(p.pix_offsets + li) : 0;
		if (MODE != MODE_LABELS && pix0 + (long long)xsteps * ysteps > p.capacity) continue;
		const og_line *ln = p.lines + li;
		const double cpx = ln->cp[0], cpy = ln->cp[1];
		const og_lin lx = og_linspace_setup(ln->cright[0], ln->cright[1], xsteps);
		const og_lin ly = og_linspace_setup(ln->up[0], ln->up[1], ysteps);
		if (MODE == MODE_LABELS)
			for (int q = lane; q <= MAX_CLASSES; q += 32) s_whist[wid][q] = 0;
		float bx0 = 3.0e38f, by0 = 3.0e38f, bx1 = -3.0e38f, by1 = -3.0e38f;
		const int cc = c0 + lane;
		const og_d2 xc = og_linspace_at(lx, min(cc, xsteps - 1));
		// rows in blocks of MAX_ROWS (one block for any OCR strip height): (ys + p0) of the block's
		// rows goes to the warp's shared-memory table   (block.py:241)
		for (int rb0 = 0; rb0 < ysteps; rb0 += MAX_ROWS) {
			const int rb_end = min(ysteps, rb0 + MAX_ROWS);
			__syncwarp();
			for (int r = rb0 + lane; r < rb_end; r += 32) {
				og_d2 yr = og_linspace_at(ly, r);
				rowbase[r - rb0] = make_double2(__dadd_rn(yr.x, cpx), __dadd_rn(yr.y, cpy));
			}
			__syncwarp();
			if (cc < xsteps) {
				uint8_t *strip = (MODE == MODE_STRIP) ? p.strips_out + pix0 + cc : nullptr;
				float2 *gout = (MODE == MODE_GRID) ? reinterpret_cast<float2 *>(p.grids_out) + pix0 + cc : nullptr;
				const float2 *__restrict__ map_lo = map;
				const float2 *__restrict__ map_hi = map + Wd;

				// one batch of RB consecutive rows: coordinates, node gather, float64 blend -> (gx, gy)
				auto coords = [&](int rbase, float (&gx)[RB], float (&gy)[RB], int (&row)[RB]) {
					int idx[RB];
					double a[RB], b[RB];
					float2 v00[RB], v01[RB], v10[RB], v11[RB];
#pragma unroll
					for (int u = 0; u < RB; ++u) {
						const int r = rbase + u;
						row[u] = r < rb_end ? r : -1;
						const int rr = min(r, rb_end - 1) - rb0;
						double rbx, rby;
						asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(rbx), "=d"(rby) : "r"(rb_saddr + rr * 16));
						// dewarped_grid = (ys + p0)[:, None] + xs[None, :]   (block.py:241)
						const double px = __dadd_rn(rbx, xc.x);
						const double py = __dadd_rn(rby, xc.y);
						// i = clip(floor(y), 0, n - 2) (SciPy find_indices); the saturating
						// conversion keeps even absurd coordinates inside the map
						const int i0 = min(max(__double2int_rd(py), 0), Hm);
						const int i1 = min(max(__double2int_rd(px), 0), Wm);
						a[u] = __dsub_rn(py, (double)i0);
						b[u] = __dsub_rn(px, (double)i1);
						idx[u] = i0 * Wd + i1;  // < 2^30 elements (dims < 32767)
						v00[u] = __ldg(map_lo + idx[u]);
						v01[u] = __ldg(map_lo + idx[u] + 1);
						v10[u] = __ldg(map_hi + idx[u]);
						v11[u] = __ldg(map_hi + idx[u] + 1);
					}
#pragma unroll
					for (int u = 0; u < RB; ++u) {
						const double wa = __dsub_rn(1.0, a[u]), wb = __dsub_rn(1.0, b[u]);
						const double w00 = __dmul_rn(wa, wb), w01 = __dmul_rn(wa, b[u]);
						const double w10 = __dmul_rn(a[u], wb), w11 = __dmul_rn(a[u], b[u]);
						const double wx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00[u].x, w00), __dmul_rn((double)v01[u].x, w01)),
							__dmul_rn((double)v10[u].x, w10)), __dmul_rn((double)v11[u].x, w11));
						const double wy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn((double)v00[u].y, w00), __dmul_rn((double)v01[u].y, w01)),
							__dmul_rn((double)v10[u].y, w10)), __dmul_rn((double)v11[u].y, w11));
						gx[u] = __double2float_rn(wx);
						gy[u] = __double2float_rn(wy);
					}
				};
				auto flush = [&](const og_taps (&t)[RB], const int (&trow)[RB]) {
#pragma unroll
					for (int u = 0; u < RB; ++u)
						if (trow[u] >= 0)  // pixels[::-1, :]  (block.py:257)
							strip[(ysteps - 1 - trow[u]) * xsteps] = (uint8_t)og_blend_taps(t[u]);
				};

				if (MODE == MODE_STRIP) {
					// two tap buffers in rotation: the taps of a batch are requested, then the previous
					// batch is blended and stored while they are in flight
					og_taps t0[RB], t1[RB];
					int r0[RB], r1[RB];
#pragma unroll
					for (int u = 0; u < RB; ++u) r1[u] = -1;
					for (int rbase = rb0; rbase < rb_end; rbase += 2 * RB) {
						float gx[RB], gy[RB];
						coords(rbase, gx, gy, r0);
						og_load_taps_batch<RB>(src, H, W, spitch, gx, gy, t0);
						flush(t1, r1);
						if (BBOX) {
#pragma unroll
							for (int u = 0; u < RB; ++u) {  // a dead slot repeats a live row: harmless for min/max
								bx0 = fminf(bx0, gx[u]); bx1 = fmaxf(bx1, gx[u]);
								by0 = fminf(by0, gy[u]); by1 = fmaxf(by1, gy[u]);
							}
						}
						coords(rbase + RB, gx, gy, r1);
						og_load_taps_batch<RB>(src, H, W, spitch, gx, gy, t1);
						flush(t0, r0);
						if (BBOX) {
#pragma unroll
							for (int u = 0; u < RB; ++u) {
								bx0 = fminf(bx0, gx[u]); bx1 = fmaxf(bx1, gx[u]);
								by0 = fminf(by0, gy[u]); by1 = fmaxf(by1, gy[u]);
							}
						}
					}
					flush(t1, r1);
				} else {
					for (int rbase = rb0; rbase < rb_end; rbase += RB) {
						float gx[RB], gy[RB];
						int row[RB];
						coords(rbase, gx, gy, row);
#pragma unroll
						for (int u = 0; u < RB; ++u) {
							if (row[u] < 0) continue;
							if (MODE == MODE_GRID) {
								gout[row[u] * xsteps] = make_float2(gx[u], gy[u]);
							} else {
								// scale_grid (batch/detect/lines.py:18-22) then cv2.remap INTER_NEAREST
								const float sx = __fmul_rn(gx[u], p.scale_x), sy = __fmul_rn(gy[u], p.scale_y);
								const int lab = og_sample_nearest_fast(src, H, W, spitch, sx, sy);
								atomicAdd(&s_whist[wid][min(lab, p.n_classes)], 1u);
							}
						}
					}
				}
			}
		}  // row block
		if (BBOX) {
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				bx0 = fminf(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
				by0 = fminf(by0, __shfl_xor_sync(0xffffffffu, by0, o));
				bx1 = fmaxf(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
				by1 = fmaxf(by1, __shfl_xor_sync(0xffffffffu, by1, o));
			}
			if (lane == 0 && bx0 <= bx1) {
				atomicMin(p.bbox_out + 4 * li + 0, (int)floorf(bx0));
				atomicMin(p.bbox_out + 4 * li + 1, (int)floorf(by0));
				atomicMax(p.bbox_out + 4 * li + 2, (int)ceilf(bx1));
				atomicMax(p.bbox_out + 4 * li + 3, (int)ceilf(by1));
			}
		}
		if (MODE == MODE_LABELS) {
			__syncwarp();
			for (int q = lane; q < p.n_classes; q += 32)
				if (s_whist[wid][q]) atomicAdd(p.counts_out + (size_t)li * p.n_classes + q, (unsigned long long)s_whist[wid][q]);
			if (lane == 0 && s_whist[wid][p.n_classes] && p.overflow_out != nullptr)
				atomicAdd(p.overflow_out + li, (unsigned long long)s_whist[wid][p.n_classes]);
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

int check_map(const float *map) {
	OG_REQUIRE((reinterpret_cast<uintptr_t>(map) & 7) == 0, OG_E_ALIGN, "map must be 8-byte aligned");
	return OG_OK;
}

int check_lines(const og_line *lines, int n, const int32_t *steps, const int32_t *chunk_offsets) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	OG_REQUIRE(n == 0 || lines != nullptr, OG_E_ARG, "lines is NULL");
	OG_REQUIRE(steps != nullptr && chunk_offsets != nullptr, OG_E_ARG, "steps / chunk_offsets is NULL");
	OG_REQUIRE((reinterpret_cast<uintptr_t>(lines) & 7) == 0, OG_E_ALIGN, "lines must be 8-byte aligned");
	return OG_OK;
}

template <int MODE, bool BBOX>
int launch_lines(const LineParams &p, cudaStream_t st) {
	if (og_first_use(40 + 2 * MODE + (BBOX ? 1 : 0))) {  // once per device and instantiation
		cudaError_t e = cudaFuncSetAttribute(og_lines_kernel<MODE, BBOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
	}
	// persistent grid: two CTAs per SM fit the staging budget
	og_lines_kernel<MODE, BBOX><<<og_num_sms() * 2, LTHREADS, SMEM_BUDGET, st>>>(p);
	return og_check_launch("og_lines_kernel");
}

template <int MODE, bool BBOX>
int launch_lines_any(const LineParams &p, const float *map, cudaStream_t st) {
	if (map == nullptr) return launch_lines<MODE, BBOX>(p, st);
	if ((MODE == MODE_STRIP || MODE == MODE_LABELS) && !BBOX && OG_LINES_SCREEN) {
		LineParams q = p;
		q.Wd = p.gw * p.cell;
		q.Hm = p.gh * p.cell - 2;
		q.Wm = q.Wd - 2;
		q.W1 = p.W - 1;
		q.H1 = p.H - 1;
		q.one_bits = 0x3F800000u;
		q.ticket = const_cast<int *>(p.chunk_offsets) + p.n + 1;
		// the ticket is zero after og_line_steps and after every completed call; it is cleared here
		// all the same, so that a plan assembled elsewhere, or a call that failed half way, cannot
		// make this one skip work
		cudaError_t e = cudaMemsetAsync(q.ticket, 0, sizeof(int), st);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
		og_strips_screened_kernel<(MODE == MODE_LABELS ? MODE_LABELS : MODE_STRIP)><<<og_num_sms() * OG_SCR_BLOCKS, STHREADS, 0, st>>>(q, reinterpret_cast<const float2 *>(map));
		return og_check_launch("og_strips_screened_kernel");
	}
	og_lines_map_kernel<MODE, BBOX><<<og_num_sms() * OG_MAP_MINBLOCKS, MTHREADS, 0, st>>>(p, reinterpret_cast<const float2 *>(map));
	return og_check_launch("og_lines_map_kernel");
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
	cudaStream_t st = (cudaStream_t)stream;
	if (n > 0 && n <= PLAN_MAX_LINES) {
		og_line_plan_kernel<<<PLAN_CTAS, 1024, 0, st>>>(reinterpret_cast<const float2 *>(grid), gh, gw, cell, lines, n,
			steps_out, pix_offsets_out, chunk_offsets_out);
		return og_check_launch("og_line_plan_kernel");
	}
	if (n > 0) {
		const int threads = 256, per_block = threads / 4;
		og_line_steps_kernel<<<(n + per_block - 1) / per_block, threads, 0, st>>>(
			reinterpret_cast<const float2 *>(grid), gh, gw, cell, lines, n, steps_out);
		rc = og_check_launch("og_line_steps_kernel");
		if (rc) return rc;
	}
	og_line_scan_kernel<<<1, 1024, 0, st>>>(steps_out, n, pix_offsets_out, chunk_offsets_out);
	return og_check_launch("og_line_scan_kernel");
}

extern "C" int og_warped_grids(const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, const int32_t *chunk_offsets,
	float *grids_out, int64_t capacity, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	rc = check_map(map);
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
	return launch_lines_any<MODE_GRID, false>(p, map, (cudaStream_t)stream);
}

int og_extract_lines_engines(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *strips_out, int64_t capacity, int32_t *bbox_out, cudaStream_t st) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	rc = check_map(map);
	if (rc) return rc;
	if (n == 0) return OG_OK;
	OG_REQUIRE(src != nullptr && H > 0 && W > 0 && src_pitch >= W, OG_E_ARG, "src invalid");
	OG_REQUIRE(H < 32767 && W < 32767, OG_E_RANGE, "source %d x %d exceeds OpenCV's 32767 limit", W, H);
	OG_REQUIRE((int64_t)H * src_pitch < ((int64_t)1 << 31), OG_E_RANGE, "source image larger than 2 GiB");
	OG_REQUIRE(pix_offsets != nullptr && strips_out != nullptr, OG_E_ARG, "pix_offsets / strips_out is NULL");
	LineParams p{};
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh; p.gw = gw; p.cell = cell;
	p.lines = lines; p.n = n;
	p.steps = steps; p.pix_offsets = pix_offsets; p.chunk_offsets = chunk_offsets;
	p.src = src; p.H = H; p.W = W; p.src_pitch = (int)src_pitch;
	p.strips_out = strips_out;
	p.bbox_out = bbox_out;
	p.capacity = capacity;
	if (bbox_out != nullptr) {
		og_bbox_init_kernel<<<(n + 255) / 256, 256, 0, st>>>(bbox_out, n);
		rc = og_check_launch("og_bbox_init_kernel");
		if (rc) return rc;
		return launch_lines_any<MODE_STRIP, true>(p, map, st);
	}
	return launch_lines_any<MODE_STRIP, false>(p, map, st);
}

extern "C" int og_sample_labels(const uint8_t *labels, int lh, int lw, int64_t labels_pitch,
	int H, int W,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, int32_t *chunk_offsets,
	int n_classes, int64_t *counts_out, int64_t *overflow_out, void *stream) {
	int rc = check_grid(grid, gh, gw, cell);
	if (rc) return rc;
	rc = check_lines(lines, n, steps, chunk_offsets);
	if (rc) return rc;
	rc = check_map(map);
	if (rc) return rc;
	OG_REQUIRE(n_classes >= 1 && n_classes <= MAX_CLASSES, OG_E_ARG, "n_classes %d outside [1, %d]", n_classes, MAX_CLASSES);
	if (n == 0) return OG_OK;
	OG_REQUIRE(labels != nullptr && lh > 0 && lw > 0 && labels_pitch >= lw, OG_E_ARG, "labels invalid");
	OG_REQUIRE(lh < 32767 && lw < 32767, OG_E_RANGE, "label map exceeds OpenCV's 32767 limit");
	OG_REQUIRE((int64_t)lh * labels_pitch < ((int64_t)1 << 31), OG_E_RANGE, "label map larger than 2 GiB");
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
	p.src = labels; p.H = lh; p.W = lw; p.src_pitch = (int)labels_pitch;
	// scale_grid: grid[:, :, 0] *= w1 / w0 with the ratio rounded to fp32
	p.scale_x = (float)((double)lw / (double)W);
	p.scale_y = (float)((double)lh / (double)H);
	p.n_classes = n_classes;
	p.counts_out = reinterpret_cast<unsigned long long *>(counts_out);
	p.overflow_out = reinterpret_cast<unsigned long long *>(overflow_out);
	return launch_lines_any<MODE_LABELS, false>(p, map, st);
}
