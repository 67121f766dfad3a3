"""
lininterp with the interface of the reference's origami/core/lingrid.py:192-301: linear
interpolation of scattered values inside the convex hull of the data points and a robust linear
extrapolation between the hull and a bounding box.  The reference evaluates two
scipy.interpolate.LinearNDInterpolator objects per query; here the two triangulations are built
once on the host (the same scipy.spatial.Delaunay call, hence the same simplices) and every query
point is located and blended on the GPU (og_tri_interp).

The construction of the extrapolation points follows lingrid.py:27-170.  The reference takes the
ray/box intersections from shapely and the corner projections from sympy; both are plain analytic
geometry and are computed here in float64 directly (shapely is not installed, so that part is not
pinned against it).
"""

import numpy as np
import scipy.spatial
import torch

from .. import ops


class _Box:
	"""lingrid.py:27-170: points on an enlarged bounding box that carry extrapolated values."""

	LEFT, TOP, BOTTOM, RIGHT = 1, 2, 3, 4

	def __init__(self, minx, miny, maxx, maxy, margin=1):
		self._minx, self._miny = minx - margin, miny - margin
		self._maxx, self._maxy = maxx + margin, maxy + margin
		self._box_size = max(maxx - minx, maxy - miny) + 2 * margin + 1
		self._points = []
		self._borders = {b: [] for b in (self.LEFT, self.TOP, self.BOTTOM, self.RIGHT)}
		self._corners = dict()

	def _intersect(self, p, direction):
		"""Where the ray p + t * direction * box_size, t in [0, 1], crosses the box outline."""
		d = direction * self._box_size
		best = None
		with np.errstate(divide="ignore", invalid="ignore"):
			for axis, bound in ((0, self._minx), (0, self._maxx), (1, self._miny), (1, self._maxy)):
				if d[axis] == 0:
					continue
				t = (bound - p[axis]) / d[axis]
				if not (0 <= t <= 1):
					continue
				q = p + t * d
				q[axis] = bound
				other = 1 - axis
				lo, hi = (self._miny, self._maxy) if axis == 0 else (self._minx, self._maxx)
				if lo <= q[other] <= hi and (best is None or t < best[0]):
					best = (t, q)
		if best is None:
			raise ValueError("point %s outside given domain" % (p,))
		x, y = float(best[1][0]), float(best[1][1])
		borders = set()
		if x == self._minx:
			borders.add(self.LEFT)
		elif x == self._maxx:
			borders.add(self.RIGHT)
		if y == self._miny:
			borders.add(self.TOP)
		elif y == self._maxy:
			borders.add(self.BOTTOM)
		return borders, x, y

	def _add(self, borders, x, y, val):
		pt = np.hstack([[x, y], val]).astype(np.float64)
		self._points.append(pt)
		for b in borders:
			self._borders[b].append(pt)

	def add_projection(self, a, b):
		a = a.astype(np.float64)
		b = b.astype(np.float64)
		v = a[:2] - b[:2]
		if np.linalg.norm(v) < 1e-2:
			return
		normal = np.array([-v[1], v[0]])
		normal /= np.linalg.norm(normal)
		self._add(*self._intersect(a[:2], normal), a[2:])
		self._add(*self._intersect(b[:2], normal), b[2:])

	def _add_corner(self, cx, cy, p1, p2):
		if p1 is None and p2 is None:
			return
		if p1 is None:
			val = p2[2:]
		elif p2 is None:
			val = p1[2:]
		else:
			# orthogonal projection of the corner onto the line p1 p2 (sympy Line.projection in the reference)
			u = p2[:2] - p1[:2]
			d_total = np.linalg.norm(u)
			if d_total == 0:
				return
			t = np.dot(np.array([cx, cy]) - p1[:2], u) / np.dot(u, u)
			p = p1[:2] + t * u
			d = np.linalg.norm(p - p1[:2])
			val = (1 - d / d_total) * p1[2:] + (d / d_total) * p2[2:]
		self._corners[(cx, cy)] = val

	def _nearest(self, border, f, axis):
		if not self._borders[border]:
			return None
		return f(self._borders[border], key=lambda pt: pt[axis])

	def add_corners(self):
		self._add_corner(self._minx, self._miny, self._nearest(self.LEFT, min, 1), self._nearest(self.TOP, min, 0))
		self._add_corner(self._maxx, self._miny, self._nearest(self.RIGHT, min, 1), self._nearest(self.TOP, max, 0))
		self._add_corner(self._maxx, self._maxy, self._nearest(self.RIGHT, max, 1), self._nearest(self.BOTTOM, max, 0))
		self._add_corner(self._minx, self._maxy, self._nearest(self.LEFT, max, 1), self._nearest(self.BOTTOM, min, 0))

	@property
	def points(self):
		for pt in self._points:
			yield pt
		for pt, val in self._corners.items():
			yield np.hstack([pt, val])


def extrapolation_points(points, values, bounds):
	"""The scattered points of the extrapolating interpolator (lingrid.py:207-262): the hull
	vertices, their projections onto the enlarged bounding box along the hull-edge normals, and the
	four corners.  Returns (extra_pts (k, 2), extra_val (k, d), is_collinear)."""
	points = np.asarray(points, dtype=np.float64)
	values = np.asarray(values, dtype=np.float64)
	if values.ndim == 1:
		values = values[:, None]
	box = _Box(*bounds)
	try:
		hull = scipy.spatial.ConvexHull(points)
		hull_idx = list(hull.vertices)
		collinear = False
	except scipy.spatial.QhullError:
		collinear = True
	if not collinear:
		extra_pts = [points[i] for i in hull_idx]
		extra_val = [values[i] for i in hull_idx]
		ring = list(zip(extra_pts, extra_val))
		for (a, va), (b, vb) in zip(ring, ring[1:] + [ring[0]]):
			box.add_projection(np.hstack([a, va]), np.hstack([b, vb]))
	else:
		extra_pts, extra_val = [], []
		pv = list(zip(points, values))
		for (pa, va), (pb, vb) in zip(pv, pv[1:]):
			a, b = np.hstack([pa, va]), np.hstack([pb, vb])
			box.add_projection(a, b)
			box.add_projection(b, a)
	box.add_corners()
	for pt in box.points:
		extra_pts.append(pt[:2])
		extra_val.append(pt[2:])
	return np.array(extra_pts, dtype=np.float64), np.array(extra_val, dtype=np.float64), collinear


class DeviceTriangulation:
	"""A Delaunay triangulation with vertex values on the GPU, plus the bucket grid og_tri_interp
	uses to find the triangle of a point."""

	def __init__(self, points, values, device, buckets_per_triangle=0.5):
		points = np.ascontiguousarray(points, dtype=np.float64)
		values = np.ascontiguousarray(values, dtype=np.float64)
		if values.ndim == 1:
			values = np.stack([values, np.zeros_like(values)], axis=1)
		if values.shape[1] != 2:
			raise ValueError("values must have one or two components")
		tri = scipy.spatial.Delaunay(points)  # qhull, as scipy.interpolate.LinearNDInterpolator does internally
		simp = tri.simplices.astype(np.int32)
		m = simp.shape[0]
		v = points[simp]  # (m, 3, 2)
		lo, hi = v.min(axis=1), v.max(axis=1)
		gmin, gmax = lo.min(axis=0), hi.max(axis=0)
		n_cells = max(1, int(m * buckets_per_triangle))
		extent = np.maximum(gmax - gmin, 1e-9)
		cell = float(np.sqrt(extent[0] * extent[1] / n_cells)) or 1.0
		gx = max(1, int(np.ceil(extent[0] / cell)))
		gy = max(1, int(np.ceil(extent[1] / cell)))
		# every triangle is listed in all buckets its bounding box touches (one bucket of slack against rounding)
		x0 = np.clip(np.floor((lo[:, 0] - gmin[0]) / cell).astype(np.int64) - 1, 0, gx - 1)
		x1 = np.clip(np.floor((hi[:, 0] - gmin[0]) / cell).astype(np.int64) + 1, 0, gx - 1)
		y0 = np.clip(np.floor((lo[:, 1] - gmin[1]) / cell).astype(np.int64) - 1, 0, gy - 1)
		y1 = np.clip(np.floor((hi[:, 1] - gmin[1]) / cell).astype(np.int64) + 1, 0, gy - 1)
		nx, ny = x1 - x0 + 1, y1 - y0 + 1
		per = nx * ny
		tri_id = np.repeat(np.arange(m, dtype=np.int64), per)
		start = np.concatenate([[0], np.cumsum(per)])[:-1]
		k = np.arange(per.sum(), dtype=np.int64) - np.repeat(start, per)
		bx = np.repeat(x0, per) + k % np.repeat(nx, per)
		by = np.repeat(y0, per) + k // np.repeat(nx, per)
		bucket = by * gx + bx
		order = np.argsort(bucket, kind="stable")
		counts = np.bincount(bucket, minlength=gx * gy)
		self.cell_start = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)).to(device)
		self.cell_tris = torch.from_numpy(tri_id[order].astype(np.int32)).to(device)
		self.gx, self.gy, self.ox, self.oy, self.cell = gx, gy, float(gmin[0]), float(gmin[1]), cell
		self.transform = torch.from_numpy(np.ascontiguousarray(tri.transform, dtype=np.float64)).to(device)
		self.simplices = torch.from_numpy(np.ascontiguousarray(simp)).to(device)
		self.values = torch.from_numpy(values).to(device)
		self.n_simplices = m

	def tables(self, eps=1e-9):
		"""struct og_tri_tables over this triangulation (the tensors must stay alive while it is used)."""
		t = ops._lib.TriTables()
		t.cell_start, t.cell_tris = self.cell_start.data_ptr(), self.cell_tris.data_ptr()
		t.transform, t.simplices, t.values = self.transform.data_ptr(), self.simplices.data_ptr(), self.values.data_ptr()
		t.ox, t.oy, t.cell, t.eps = self.ox, self.oy, self.cell, float(eps)
		t.gx, t.gy = self.gx, self.gy
		return t

	def interp(self, pts_dev, clip_box, out=None, only_nan=False, eps=1e-9):
		if out is None:
			out = torch.empty_like(pts_dev)
		box = np.ascontiguousarray(clip_box, dtype=np.float64)
		ops._call("og_tri_interp", pts_dev, ([out, self.transform], lambda st: (
			pts_dev.data_ptr(), pts_dev.shape[0], box.ctypes.data,
			self.cell_start.data_ptr(), self.cell_tris.data_ptr(), self.gx, self.gy, self.ox, self.oy, self.cell,
			self.transform.data_ptr(), self.simplices.data_ptr(), self.values.data_ptr(), float(eps),
			out.data_ptr(), 1 if only_nan else 0, st)))
		return out


class Interpolator:
	"""lingrid.py:173-189: clip to the bounds, interpolate inside the hull, extrapolate elsewhere."""

	def __init__(self, inter, extra, bounds, squeeze, device):
		self._inter, self._extra, self._bounds, self._squeeze, self._device = inter, extra, bounds, squeeze, device

	def field_struct(self):
		"""struct og_field for the kernels that evaluate this (scalar) interpolant on the device."""
		f = ops._lib.Field()
		f.extra = self._extra.tables()
		if self._inter is not None:
			f.inter = self._inter.tables()
		f.has_inter = 0 if self._inter is None else 1
		f.constant = 0
		f.phi0 = 0.0
		for k in range(4):
			f.bounds[k] = float(self._bounds[k])
		return f

	@property
	def device(self):
		return self._device

	def device_call(self, pts_dev):
		"""(n, 2) float64 CUDA tensor -> (n, 2) float64 CUDA tensor."""
		if self._inter is None:
			return self._extra.interp(pts_dev, self._bounds)
		out = self._inter.interp(pts_dev, self._bounds)
		return self._extra.interp(pts_dev, self._bounds, out=out, only_nan=True)

	def __call__(self, pts):
		pts = np.array(pts, dtype=np.float64)
		if pts.ndim == 1:
			pts = pts[np.newaxis, :]
		dev = torch.from_numpy(np.ascontiguousarray(pts)).to(self._device)
		res = self.device_call(dev).cpu().numpy()
		return res[:, 0] if self._squeeze else res


def lininterp(points, values, bounds, device=None):
	"""lingrid.py:300-301.  values: (n,) scalars or (n, 2) vectors."""
	points = np.asarray(points, dtype=np.float64)
	values_a = np.asarray(values, dtype=np.float64)
	if len(values_a) < 2:
		raise ValueError("not enough values given.")
	if device is None:
		if not torch.cuda.is_available():
			raise ops._lib.OrigamiB200Error("origami_b200 needs a CUDA device (there is no CPU fallback)")
		device = torch.device("cuda", torch.cuda.current_device())
	squeeze = values_a.ndim == 1
	extra_pts, extra_val, collinear = extrapolation_points(points, values_a, bounds)
	if squeeze:
		extra_val = extra_val[:, 0]
	extra = DeviceTriangulation(extra_pts, extra_val, device)
	inter = None if collinear else DeviceTriangulation(points, values_a, device)
	return Interpolator(inter, extra, tuple(float(b) for b in bounds), squeeze, device)
