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

struct TriParams {
	const double *pts;       // (n, 2) query points (x, y)
	long long n;
	double minx, miny, maxx, maxy;  // clip box applied to the queries (lingrid.py:196-197)
	const int32_t *cell_start;      // (gy * gx + 1) offsets into cell_tris
	const int32_t *cell_tris;       // triangle indices per bucket
	int gx, gy;
	double ox, oy, inv_cell;        // bucket grid origin and 1 / bucket size
	const double *transform;        // (m, 3, 2): rows 0-1 = inverse edge matrix, row 2 = reference vertex (scipy Delaunay.transform)
	const int32_t *simplices;       // (m, 3) vertex indices
	const double *values;           // (n_vertices, 2)
	double eps;                     // barycentric tolerance
	double *out;                    // (n, 2)
	int only_nan;                   // fill only the entries that hold NaN (second triangulation of lingrid's Interpolator)
};

__global__ void __launch_bounds__(256) og_tri_interp_kernel(const __grid_constant__ TriParams p) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= p.n) return;
	if (p.only_nan) {
		const double o = p.out[2 * i];
		if (o == o) return;
	}
	double x = p.pts[2 * i], y = p.pts[2 * i + 1];
	x = fmin(fmax(x, p.minx), p.maxx);  // np.clip
	y = fmin(fmax(y, p.miny), p.maxy);
	const double nan = __longlong_as_double(0x7ff8000000000000ll);
	double rx = nan, ry = nan;
	if (x == x && y == y) {
		const int cx = min(max((int)floor((x - p.ox) * p.inv_cell), 0), p.gx - 1);
		const int cy = min(max((int)floor((y - p.oy) * p.inv_cell), 0), p.gy - 1);
		const int beg = p.cell_start[cy * p.gx + cx], end = p.cell_start[cy * p.gx + cx + 1];
		double best = -1.0e300;
		int best_t = -1;
		double bc0 = 0, bc1 = 0;
		for (int k = beg; k < end; ++k) {
			const int t = p.cell_tris[k];
			const double *T = p.transform + (size_t)t * 6;
			const double dx = x - T[4], dy = y - T[5];
			const double c0 = T[0] * dx + T[1] * dy;
			const double c1 = T[2] * dx + T[3] * dy;
			const double c2 = 1.0 - c0 - c1;
			const double m = fmin(c0, fmin(c1, c2));  // NaN transforms (degenerate simplices) never win
			if (m > best) {
				best = m;
				best_t = t;
				bc0 = c0;
				bc1 = c1;
			}
		}
		if (best_t >= 0 && best >= -p.eps) {
			const int32_t *s = p.simplices + (size_t)best_t * 3;
			const double *v0 = p.values + 2 * (size_t)s[0], *v1 = p.values + 2 * (size_t)s[1], *v2 = p.values + 2 * (size_t)s[2];
			const double c2 = 1.0 - bc0 - bc1;
			rx = bc0 * v0[0] + bc1 * v1[0] + c2 * v2[0];
			ry = bc0 * v0[1] + bc1 * v1[1] + c2 * v2[1];
		}
	}
	p.out[2 * i] = rx;
	p.out[2 * i + 1] = ry;
}

}  // namespace

extern "C" int og_tri_interp(const double *pts, int64_t n, const double *clip_box,
	const int32_t *cell_start, const int32_t *cell_tris, int gx, int gy, double ox, double oy, double cell,
	const double *transform, const int32_t *simplices, const double *values, double eps,
	double *out, int only_nan, void *stream) {
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	if (n == 0) return OG_OK;
	OG_REQUIRE(pts != nullptr && out != nullptr && clip_box != nullptr, OG_E_ARG, "pts / out / clip_box is NULL");
	OG_REQUIRE(cell_start != nullptr && cell_tris != nullptr && transform != nullptr && simplices != nullptr && values != nullptr,
		OG_E_ARG, "triangulation table is NULL");
	OG_REQUIRE(gx > 0 && gy > 0 && cell > 0.0, OG_E_ARG, "bucket grid %d x %d / size invalid", gx, gy);
	TriParams p{};
	p.pts = pts; p.n = n;
	p.minx = clip_box[0]; p.miny = clip_box[1]; p.maxx = clip_box[2]; p.maxy = clip_box[3];  // HOST array of 4 doubles
	p.cell_start = cell_start; p.cell_tris = cell_tris;
	p.gx = gx; p.gy = gy; p.ox = ox; p.oy = oy; p.inv_cell = 1.0 / cell;
	p.transform = transform; p.simplices = simplices; p.values = values;
	p.eps = eps;
	p.out = out; p.only_nan = only_nan;
	og_tri_interp_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
	return og_check_launch("og_tri_interp_kernel");
}
