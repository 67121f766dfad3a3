// Scattered linear interpolation on a triangulation: the warped ->
// dewarped point transform of Grid.transformer (origami/core/dewarp.py:143-161, :555-559), which
// origami/batch/detect/dewarp.py:16-39 applies to every vertex of every contour polygon and
// Line.dewarped_height (origami/core/block.py:326-333) to the two ends of a line's `up` vector.
//
// The reference evaluates scipy.interpolate.LinearNDInterpolator over the warped lattice nodes
// (origami/core/lingrid.py:266-275): Delaunay triangulation by qhull, point location by a walk,
// barycentric blend of the three vertex values.  The triangulation is built on the host by the
// same SciPy call (so the diagonals are the oracle's) and uploaded once per page; this kernel does
// the per-point work for all vertices of a page in one launch: a uniform bucket grid over the
// triangles' bounding boxes replaces the walk, the barycentric coordinates come from SciPy's own
// per-simplex affine transform (Delaunay.transform), the blend is float64.
#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

// value of the piecewise-linear interpolant at (x, y); false (and NaN) where no triangle holds the point
__device__ __forceinline__ bool tri_eval(const og_tri_tables &t, double x, double y, double &rx, double &ry) {
	const double nan = __longlong_as_double(0x7ff8000000000000ll);
	rx = ry = nan;
	if (!(x == x && y == y)) return false;
	const double inv_cell = 1.0 / t.cell;
	const int cx = min(max((int)floor((x - t.ox) * inv_cell), 0), t.gx - 1);
	const int cy = min(max((int)floor((y - t.oy) * inv_cell), 0), t.gy - 1);
	const int beg = t.cell_start[cy * t.gx + cx], end = t.cell_start[cy * t.gx + cx + 1];
	double best = -1.0e300;
	int best_t = -1;
	double bc0 = 0, bc1 = 0;
	for (int k = beg; k < end; ++k) {
		const int ti = t.cell_tris[k];
		const double *T = t.transform + (size_t)ti * 6;
		const double dx = x - T[4], dy = y - T[5];
		const double c0 = T[0] * dx + T[1] * dy;
		const double c1 = T[2] * dx + T[3] * dy;
		const double c2 = 1.0 - c0 - c1;
		const double m = fmin(c0, fmin(c1, c2));  // NaN transforms (degenerate simplices) never win
		if (m > best) {
			best = m;
			best_t = ti;
			bc0 = c0;
			bc1 = c1;
		}
	}
	if (best_t < 0 || best < -t.eps) return false;
	const int32_t *s = t.simplices + (size_t)best_t * 3;
	const double *v0 = t.values + 2 * (size_t)s[0], *v1 = t.values + 2 * (size_t)s[1], *v2 = t.values + 2 * (size_t)s[2];
	const double c2 = 1.0 - bc0 - bc1;
	rx = bc0 * v0[0] + bc1 * v1[0] + c2 * v2[0];
	ry = bc0 * v0[1] + bc1 * v1[1] + c2 * v2[1];
	return true;
}

__global__ void __launch_bounds__(256) og_tri_interp_kernel(const __grid_constant__ og_tri_tables t, const double *__restrict__ pts, long long n,
	double minx, double miny, double maxx, double maxy, double *__restrict__ out, int only_nan) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	if (only_nan) {
		const double o = out[2 * i];
		if (o == o) return;
	}
	const double x = fmin(fmax(pts[2 * i], minx), maxx);  // np.clip (lingrid.py:196-197)
	const double y = fmin(fmax(pts[2 * i + 1], miny), maxy);
	double rx, ry;
	tri_eval(t, x, y, rx, ry);
	out[2 * i] = rx;
	out[2 * i + 1] = ry;
}

// lingrid's Interpolator for a scalar field (component 0 of the values): clip, interpolate inside the
// hull, extrapolate elsewhere; a field built from fewer than three samples is the constant phi0
// (origami/core/dewarp.py:112-116)
__device__ __forceinline__ double field_angle(const og_field &f, double x, double y) {
	if (f.constant) return f.phi0;
	x = fmin(fmax(x, f.bounds[0]), f.bounds[2]);
	y = fmin(fmax(y, f.bounds[1]), f.bounds[3]);
	double a, unused;
	if (f.has_inter && tri_eval(f.inter, x, y, a, unused)) return a;
	tri_eval(f.extra, x, y, a, unused);
	return a;
}

// Field.get (dewarp.py:118-122): np.cos / np.sin of the angles evaluated as float32
__device__ __forceinline__ float2 field_dir(const og_field &f, float px, float py) {
	const float a = (float)field_angle(f, (double)px, (double)py);
	return make_float2(cosf(a), sinf(a));
}

// Euler integration through a direction field (GridFactory.grid_h, extend_border_*, Field.estimate_extent,
// dewarp.py:123-141, 341-439): P points advance n_steps times by `step` along the field; out[k] holds the
// positions BEFORE step k (out[0] = the start points), all float32 as in the reference.
__global__ void __launch_bounds__(128) og_field_march_kernel(const __grid_constant__ og_field f, const float2 *__restrict__ start, int P,
	int n_steps, float step, float2 *__restrict__ out) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P) return;
	float2 p = start[i];
	for (int k = 0; k < n_steps; ++k) {
		out[(size_t)k * P + i] = p;
		const float2 d = field_dir(f, p.x, p.y);
		p.x = __fadd_rn(p.x, __fmul_rn(d.x, step));  // pts += field(pts) * step
		p.y = __fadd_rn(p.y, __fmul_rn(d.y, step));
	}
}

// GridFactory.grid_hv's column marching (dewarp.py:441-470 with BatchIntersections :176-263): a thread owns
// a column of the lattice; from its node on h-line gy it casts a ray of length max_step_len along the
// vertical field and moves to the nearest crossing with the polyline of h-line gy + 1 (the next rows are
// tried when that one is missed).  grid_h holds the h-lines with their end points already pushed outwards
// (make_row_slice does that in place in the reference).
__global__ void __launch_bounds__(64) og_grid_vlines_kernel(const __grid_constant__ og_field fv, const float2 *__restrict__ grid_h, int gh, int gw,
	int col0, int ncols, float max_step_len, float2 *__restrict__ grid_hv, int *__restrict__ status) {
	const int c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= ncols) return;
	const int col = col0 + c;
	float2 p0 = grid_h[col];
	for (int gy = 0; gy < gh - 1; ++gy) {
		grid_hv[(size_t)gy * gw + col] = p0;
		const float2 d = field_dir(fv, p0.x, p0.y);
		const float2 p1 = make_float2(__fadd_rn(p0.x, __fmul_rn(d.x, max_step_len)), __fadd_rn(p0.y, __fmul_rn(d.y, max_step_len)));
		const double ax = p0.x, ay = p0.y, rx = (double)p1.x - ax, ry = (double)p1.y - ay;
		bool found = false;
		double bx = 0, by = 0;
		for (int row = gy + 1; row < gh && !found; ++row) {  // `dy += 1  # retry on next row`
			double best = 1.0e300;
			const float2 *line = grid_h + (size_t)row * gw;
			for (int j = 0; j + 1 < gw; ++j) {
				const double qx = line[j].x, qy = line[j].y, sx = (double)line[j + 1].x - qx, sy = (double)line[j + 1].y - qy;
				const double den = rx * sy - ry * sx;
				if (den == 0.0) continue;
				const double t = ((qx - ax) * sy - (qy - ay) * sx) / den;  // along the ray
				const double u = ((qx - ax) * ry - (qy - ay) * rx) / den;  // along the segment
				if (t >= 0.0 && t <= 1.0 && u >= 0.0 && u <= 1.0) {
					const double ix = ax + t * rx, iy = ay + t * ry;
					const double dist = (ix - ax) * (ix - ax) + (iy - ay) * (iy - ay);
					if (dist < best) {
						best = dist;
						bx = ix;
						by = iy;
						found = true;
					}
				}
			}
		}
		if (!found) {
			atomicExch(status, 1 + col);  // "failed to find intersection" (dewarp.py:244-246)
			bx = p1.x;
			by = p1.y;
		}
		p0 = make_float2((float)bx, (float)by);
	}
	grid_hv[(size_t)(gh - 1) * gw + col] = p0;
}

}  // namespace

static int check_tables(const og_tri_tables *t, const char *what) {
	OG_REQUIRE(t->cell_start != nullptr && t->cell_tris != nullptr && t->transform != nullptr && t->simplices != nullptr && t->values != nullptr,
		OG_E_ARG, "%s: triangulation table is NULL", what);
	OG_REQUIRE(t->gx > 0 && t->gy > 0 && t->cell > 0.0, OG_E_ARG, "%s: bucket grid %d x %d / size invalid", what, t->gx, t->gy);
	return OG_OK;
}

static int check_field(const og_field *f) {
	OG_REQUIRE(f != nullptr, OG_E_ARG, "field is NULL");
	if (f->constant) return OG_OK;
	int rc = check_tables(&f->extra, "field.extra");
	if (rc) return rc;
	return f->has_inter ? check_tables(&f->inter, "field.inter") : OG_OK;
}

extern "C" int og_tri_interp(const double *pts, int64_t n, const double *clip_box,
	const int32_t *cell_start, const int32_t *cell_tris, int gx, int gy, double ox, double oy, double cell,
	const double *transform, const int32_t *simplices, const double *values, double eps,
	double *out, int only_nan, void *stream) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	if (n == 0) return OG_OK;
	OG_REQUIRE(pts != nullptr && out != nullptr && clip_box != nullptr, OG_E_ARG, "pts / out / clip_box is NULL");
	og_tri_tables t{};
	t.cell_start = cell_start; t.cell_tris = cell_tris;
	t.gx = gx; t.gy = gy; t.ox = ox; t.oy = oy; t.cell = cell;
	t.transform = transform; t.simplices = simplices; t.values = values;
	t.eps = eps;
	int rc = check_tables(&t, "og_tri_interp");
	if (rc) return rc;
	og_tri_interp_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(t, pts, n,
		clip_box[0], clip_box[1], clip_box[2], clip_box[3], out, only_nan);  // clip_box: HOST array of 4 doubles
	return og_check_launch("og_tri_interp_kernel");
}

extern "C" int og_field_march(const og_field *field, const float *start_xy, int n_points, int n_steps, float step,
	float *out_xy, void *stream) {
	int rc = check_field(field);
	if (rc) return rc;
	OG_REQUIRE(n_points >= 0 && n_steps >= 0, OG_E_ARG, "negative count");
	if (n_points == 0 || n_steps == 0) return OG_OK;
	OG_REQUIRE(start_xy != nullptr && out_xy != nullptr, OG_E_ARG, "NULL pointer");
	og_field_march_kernel<<<(n_points + 127) / 128, 128, 0, (cudaStream_t)stream>>>(*field, reinterpret_cast<const float2 *>(start_xy),
		n_points, n_steps, step, reinterpret_cast<float2 *>(out_xy));
	return og_check_launch("og_field_march_kernel");
}

extern "C" int og_grid_vlines(const og_field *field_v, const float *grid_h, int gh, int gw, int col0, int ncols,
	float max_step_len, float *grid_hv_out, int32_t *status_out, void *stream) {
	int rc = check_field(field_v);
	if (rc) return rc;
	OG_REQUIRE(grid_h != nullptr && grid_hv_out != nullptr && status_out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(gh >= 2 && gw >= 2 && col0 >= 0 && ncols >= 0 && col0 + ncols <= gw, OG_E_ARG, "grid (%d, %d) / column slice invalid", gh, gw);
	if (ncols == 0) return OG_OK;
	og_grid_vlines_kernel<<<(ncols + 63) / 64, 64, 0, (cudaStream_t)stream>>>(*field_v, reinterpret_cast<const float2 *>(grid_h), gh, gw,
		col0, ncols, max_step_len, reinterpret_cast<float2 *>(grid_hv_out), status_out);
	return og_check_launch("og_grid_vlines_kernel");
}

// ---------------------------------------------------------------------------------------
// Polygon masks (A11): origami/core/mask.py:8-48 fills the polygon (vertices rounded to integers)
// with cairo and keeps the pixels whose red channel is non-zero, i.e. every pixel the polygon
// covers by at least half an 8-bit alpha step.  cairo is not available, so the coverage is
// computed exactly instead: the area of polygon /\ pixel square by Green's theorem, edge by edge,
// each edge's clamped abscissa integrated over the pixel's row in closed form (float64).  A thread
// owns a pixel and walks all edges in order, so the result does not depend on scheduling.
// ---------------------------------------------------------------------------------------
namespace {

// integral over y in [ya, yb] of clamp(x(y) - px, 0, 1), x linear with x(ya) = xa, x(yb) = xb
__device__ __forceinline__ double clamped_strip_area(double ya, double yb, double xa, double xb, double px) {
	const double fa = xa - px, fb = xb - px, len = yb - ya;
	if (len <= 0.0) return 0.0;
	if (fa <= 0.0 && fb <= 0.0) return 0.0;
	if (fa >= 1.0 && fb >= 1.0) return len;
	// breakpoints where the edge crosses the left (f = 0) and the right (f = 1) side of the pixel column
	const double df = fb - fa;
	double t0 = 0.0, t1 = 1.0;  // parameters of the crossings, sorted below
	double area = 0.0, tprev = 0.0, fprev = fmin(fmax(fa, 0.0), 1.0);
	double ts[2];
	int nt = 0;
	if (df != 0.0) {
		t0 = (0.0 - fa) / df;
		t1 = (1.0 - fa) / df;
		const double lo = fmin(t0, t1), hi = fmax(t0, t1);
		if (lo > 0.0 && lo < 1.0) ts[nt++] = lo;
		if (hi > 0.0 && hi < 1.0) ts[nt++] = hi;
	}
	for (int k = 0; k <= nt; ++k) {
		const double t = k < nt ? ts[k] : 1.0;
		const double f = fmin(fmax(fa + df * t, 0.0), 1.0);
		area += 0.5 * (fprev + f) * (t - tprev);
		tprev = t;
		fprev = f;
	}
	return area * len;
}

__global__ void __launch_bounds__(256) og_polygon_mask_kernel(const double *__restrict__ verts, int k, int w, int h,
	uint8_t *__restrict__ mask, int64_t mask_pitch, uint8_t *__restrict__ img, int64_t img_pitch, int background) {
	const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
	if (x >= w || y >= h) return;
	const double px = (double)x, py = (double)y;
	double acc = 0.0;
	for (int e = 0; e < k; ++e) {
		const int e1 = e + 1 == k ? 0 : e + 1;
		const double x0 = verts[2 * e], y0 = verts[2 * e + 1], x1 = verts[2 * e1], y1 = verts[2 * e1 + 1];
		if (y0 == y1) continue;
		const double ylo = fmin(y0, y1), yhi = fmax(y0, y1);
		const double ya = fmax(ylo, py), yb = fmin(yhi, py + 1.0);
		if (ya >= yb) continue;
		const double m = (x1 - x0) / (y1 - y0);
		const double xa = x0 + (ya - y0) * m, xb = x0 + (yb - y0) * m;
		const double a = clamped_strip_area(ya, yb, xa, xb, px);
		acc += y1 > y0 ? a : -a;
	}
	const bool inside = fabs(acc) >= 0.5 / 255.0;  // alpha = round(255 * coverage) > 0
	if (mask != nullptr) mask[(int64_t)y * mask_pitch + x] = inside ? 255 : 0;
	if (img != nullptr && !inside) img[(int64_t)y * img_pitch + x] = (uint8_t)background;
}

}  // namespace

extern "C" int og_polygon_mask(const double *verts_xy, int k, int w, int h,
	uint8_t *mask_out, int64_t mask_pitch, uint8_t *img_inout, int64_t img_pitch, int background, void *stream) {
	OG_REQUIRE(w > 0 && h > 0, OG_E_ARG, "mask size %d x %d invalid", w, h);
	OG_REQUIRE(k >= 0 && (k == 0 || verts_xy != nullptr), OG_E_ARG, "vertices missing");
	OG_REQUIRE(mask_out != nullptr || img_inout != nullptr, OG_E_ARG, "neither mask_out nor img_inout given");
	OG_REQUIRE((mask_out == nullptr || mask_pitch >= w) && (img_inout == nullptr || img_pitch >= w), OG_E_ARG, "pitch smaller than row");
	dim3 grid((w + 255) / 256, h);
	OG_REQUIRE(h <= 65535, OG_E_RANGE, "mask taller than 65535 rows");
	og_polygon_mask_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(verts_xy, k, w, h, mask_out, mask_pitch, img_inout, img_pitch, background);
	return og_check_launch("og_polygon_mask_kernel");
}

// ---------------------------------------------------------------------------------------
// SplitDetector's column statistic (origami/batch/detect/layout.py:
// 915-944): scipy.signal.periodogram(pixels, axis=0) -- per image column the one-sided power
// spectral density of the mean-free column (boxcar window, density scaling) -- followed by
// np.quantile(dens, q, axis=0).  One CTA per column: the column and a table of the N-th roots of
// unity sit in shared memory, every thread evaluates a few frequencies of the DFT directly in
// float64 (N is a region height, not a power of two; N^2 / 2 terms per column is small work for a
// GPU and has no FFT round-off pattern to match), the densities are sorted in shared memory and
// the quantile is read off with NumPy's linear interpolation rule.
// ---------------------------------------------------------------------------------------
namespace {

// One series of h samples (element j at pixels[first + j * stride]) -> *out; the CTA's dynamic shared
// memory holds h roots of unity, h samples and nsort (power of two >= h / 2 + 1) densities.
__device__ __forceinline__ void psd_quantile_series(const void *__restrict__ pixels, int is_u8, int64_t first, int64_t stride,
	int h, double q, float *__restrict__ out, int nsort, unsigned char *smem_raw) {
	double2 *tw = reinterpret_cast<double2 *>(smem_raw);                   // h roots of unity
	float *col = reinterpret_cast<float *>(smem_raw + (size_t)h * 16);     // h samples
	float *dens = col + h;                                                  // nsort (power of two) densities
	__shared__ double s_part[8];
	__shared__ double s_mean;
	const int nf = h / 2 + 1;
	for (int j = threadIdx.x; j < h; j += blockDim.x) {
		double sn, cs;
		sincospi(-2.0 * (double)j / (double)h, &sn, &cs);
		tw[j] = make_double2(cs, sn);
		float v;
		if (is_u8) v = __fdiv_rn((float)reinterpret_cast<const uint8_t *>(pixels)[first + (int64_t)j * stride], 255.0f);  // pixels.astype(float32) / 255.0
		else v = reinterpret_cast<const float *>(pixels)[first + (int64_t)j * stride];
		col[j] = v;
	}
	__syncthreads();
	// detrend="constant": subtract the column mean
	double part = 0.0;
	for (int j = threadIdx.x; j < h; j += blockDim.x) part += (double)col[j];
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
	if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = part;
	__syncthreads();
	if (threadIdx.x == 0) {
		double t = 0.0;
		for (int k = 0; k < (int)(blockDim.x >> 5); ++k) t += s_part[k];
		s_mean = t / (double)h;
	}
	__syncthreads();
	const float mean = (float)s_mean;
	for (int j = threadIdx.x; j < h; j += blockDim.x) col[j] = __fsub_rn(col[j], mean);
	__syncthreads();
	for (int k = threadIdx.x; k < nsort; k += blockDim.x) {
		float d = __int_as_float(0x7f800000);  // padding sorts to the end
		if (k < nf) {
			double re = 0.0, im = 0.0;
			int idx = 0;
			for (int n = 0; n < h; ++n) {
				const double2 t = tw[idx];
				const double v = (double)col[n];
				re += v * t.x;
				im += v * t.y;
				idx += k;
				if (idx >= h) idx -= h;
			}
			double p = (re * re + im * im) / (double)h;
			const bool doubled = k > 0 && ((h & 1) ? true : k < h / 2);
			if (doubled) p *= 2.0;
			d = (float)p;
		}
		dens[k] = d;
	}
	__syncthreads();
	// bitonic sort of nsort values
	for (int size = 2; size <= nsort; size <<= 1) {
		for (int stride = size >> 1; stride > 0; stride >>= 1) {
			for (int i = threadIdx.x; i < nsort; i += blockDim.x) {
				const int j = i ^ stride;
				if (j > i) {
					const bool up = (i & size) == 0;
					const float a = dens[i], b = dens[j];
					if ((a > b) == up) {
						dens[i] = b;
						dens[j] = a;
					}
				}
			}
			__syncthreads();
		}
	}
	if (threadIdx.x == 0) {
		// np.quantile, method "linear": virtual index q * (n - 1), lerp with NumPy's _lerp
		const double pos = q * (double)(nf - 1);
		int lo = (int)floor(pos);
		lo = min(max(lo, 0), nf - 1);
		const int hi = min(lo + 1, nf - 1);
		const float t = (float)(pos - (double)lo);
		const float a = dens[lo], b = dens[hi];
		const float diff = __fsub_rn(b, a);
		float r = __fadd_rn(a, __fmul_rn(diff, t));
		if (t >= 0.5f) r = __fsub_rn(b, __fmul_rn(diff, __fsub_rn(1.f, t)));
		if (t == 1.f || !(diff == diff)) r = b;
		*out = r;
	}
}

__global__ void __launch_bounds__(256) og_column_psd_quantile_kernel(const void *__restrict__ pixels, int is_u8, int64_t pitch_elems,
	int h, int w, double q, float *__restrict__ out, int nsort) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	psd_quantile_series(pixels, is_u8, blockIdx.x, pitch_elems, h, q, out + blockIdx.x, nsort, smem_raw);
}

// All regions of a page in one launch: CTA b works on series b - offsets[r] of region r (offsets = prefix
// sums of the regions' series counts).  A region is (x0, y0, w, h, transposed): h samples down each of
// its w columns, or -- transposed, FixSpillOverV's crop.transpose() -- w samples along each of its h rows.
__global__ void __launch_bounds__(256) og_region_psd_quantile_kernel(const void *__restrict__ pixels, int is_u8, int64_t pitch_elems,
	const int32_t *__restrict__ regions, const int32_t *__restrict__ offsets, int m, double q, float *__restrict__ out) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int b = blockIdx.x;
	int lo = 0, hi = m;
	while (hi - lo > 1) {
		const int mid = (lo + hi) >> 1;
		if (__ldg(offsets + mid) <= b) lo = mid; else hi = mid;
	}
	const int c = b - __ldg(offsets + lo);
	const int x0 = regions[5 * lo], y0 = regions[5 * lo + 1], w = regions[5 * lo + 2], h = regions[5 * lo + 3];
	const bool tr = regions[5 * lo + 4] != 0;
	const int len = tr ? w : h;
	if (c >= (tr ? h : w) || len < 1) return;
	int nsort = 1;
	while (nsort < len / 2 + 1) nsort <<= 1;
	const int64_t first = tr ? (int64_t)(y0 + c) * pitch_elems + x0 : (int64_t)y0 * pitch_elems + x0 + c;
	psd_quantile_series(pixels, is_u8, first, tr ? (int64_t)1 : pitch_elems, len, q, out + b, nsort, smem_raw);
}

}  // namespace

extern "C" int og_column_psd_quantile(const void *pixels, int is_u8, int h, int w, int64_t pitch_bytes,
	double quantile, float *out, void *stream) {
	OG_REQUIRE(pixels != nullptr && out != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(h >= 1 && w >= 1, OG_E_ARG, "crop %d x %d invalid", w, h);
	OG_REQUIRE(h <= 8192, OG_E_RANGE, "crop of %d rows exceeds the 8192-row limit of the shared-memory DFT", h);
	OG_REQUIRE(quantile >= 0.0 && quantile <= 1.0, OG_E_ARG, "quantile %g outside [0, 1]", quantile);
	const int elem = is_u8 ? 1 : 4;
	OG_REQUIRE(pitch_bytes >= (int64_t)w * elem && pitch_bytes % elem == 0, OG_E_ARG, "pitch invalid");
	int nsort = 1;
	while (nsort < h / 2 + 1) nsort <<= 1;
	const size_t smem = (size_t)h * 16 + (size_t)h * 4 + (size_t)nsort * 4;
	if (og_first_use(57)) {
		cudaError_t e = cudaFuncSetAttribute(og_column_psd_quantile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
	}
	og_column_psd_quantile_kernel<<<w, 256, smem, (cudaStream_t)stream>>>(pixels, is_u8, pitch_bytes / elem, h, w, quantile, out, nsort);
	return og_check_launch("og_column_psd_quantile_kernel");
}

extern "C" int og_region_psd_quantile(const void *pixels, int is_u8, int H, int W, int64_t pitch_bytes,
	const int32_t *regions, const int32_t *offsets, int m, int total, int max_len, double quantile, float *out, void *stream) {
	OG_REQUIRE(m >= 0 && total >= 0, OG_E_ARG, "m or total < 0");
	if (m == 0 || total == 0) return OG_OK;
	OG_REQUIRE(pixels != nullptr && out != nullptr && regions != nullptr && offsets != nullptr, OG_E_ARG, "NULL pointer");
	OG_REQUIRE(H >= 1 && W >= 1, OG_E_ARG, "image %d x %d invalid", W, H);
	OG_REQUIRE(max_len >= 1 && max_len <= 8192, OG_E_RANGE, "series of %d samples exceed the 8192-sample limit of the shared-memory DFT", max_len);
	OG_REQUIRE(quantile >= 0.0 && quantile <= 1.0, OG_E_ARG, "quantile %g outside [0, 1]", quantile);
	const int elem = is_u8 ? 1 : 4;
	OG_REQUIRE(pitch_bytes >= (int64_t)W * elem && pitch_bytes % elem == 0, OG_E_ARG, "pitch invalid");
	int nsort = 1;
	while (nsort < max_len / 2 + 1) nsort <<= 1;
	const size_t smem = (size_t)max_len * 16 + (size_t)max_len * 4 + (size_t)nsort * 4;
	if (og_first_use(58)) {
		cudaError_t e = cudaFuncSetAttribute(og_region_psd_quantile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
	}
	og_region_psd_quantile_kernel<<<total, 256, smem, (cudaStream_t)stream>>>(pixels, is_u8, pitch_bytes / elem, regions, offsets, m, quantile, out);
	return og_check_launch("og_region_psd_quantile_kernel");
}
