"""
Samples, Field and GridFactory with the interface of the reference's origami/core/dewarp.py:34-141
and :292-493 (Schneider / Block / Rojas, "Robust document warping with interpolated vector fields",
ICDAR 2007): flow samples (x, y, angle) -> two interpolated angle fields -> h-lines by Euler
integration -> v-lines by marching from h-line to h-line along the vertical field -> border
extension -> the sparse (gh, gw, 2) grid that `Grid` wraps.

The reference evaluates the fields through scipy interpolators inside Python loops (one call per
lattice column / row) and intersects rays with shapely STRtrees on two threads.  Here the fields
live on the GPU as triangulations (core/lingrid.py) and every loop is one kernel launch:
og_field_march integrates all rows (or columns) of a step sequence at once, og_grid_vlines marches
all columns, each thread finding its ray's nearest crossing with the next h-line itself.
Floating point: positions are float32 as in the reference; cos / sin of the angles are CUDA's
float32 functions (NumPy's differ in the last ulp), the crossings plain float64 segment
intersections (GEOS in the reference) -- agreement to about 1e-3 px over a page, not bit parity.
"""

import io
import json
import math
from functools import cached_property

import numpy as np
import torch

from .. import _lib, ops
from .lingrid import lininterp
from .math import Geometry


class Samples:
	"""Scattered (point, angle) samples of one flow direction and their flow.zip encoding (reference
	dewarp.py:34-105): NAME.npy holds float64 rows (x, y, angle) -- an empty set is written as a (3, 0)
	array --, NAME.json the page size and a format version."""

	def __init__(self, geometry):
		self._geometry = geometry
		self._points, self._values = [], []

	def __len__(self):
		return len(self._points)

	@staticmethod
	def open(zf, name):
		meta = json.loads(zf.read(name + ".json"))
		table = np.load(io.BytesIO(zf.read(name + ".npy")), allow_pickle=False)
		loaded = Samples(Geometry(*meta["size"]))
		if table.size:
			loaded.extend(table[:, :2], table[:, 2])
		return loaded

	def save(self, zf, name):
		if len(self):
			table = np.column_stack([np.asarray(self._points, dtype=np.float64).reshape(-1, 2), np.asarray(self._values, dtype=np.float64)])
		else:
			table = np.empty((3, 0), dtype=np.float64)
		buf = io.BytesIO()
		np.save(buf, table, allow_pickle=False)
		zf.writestr(name + ".npy", buf.getvalue())
		zf.writestr(name + ".json", json.dumps({"version": 1, "size": self._geometry.size}))

	geometry = property(lambda self: self._geometry)
	points = property(lambda self: self._points)
	values = property(lambda self: self._values)

	def append(self, point, value):
		self.extend([point], [value])

	def extend(self, points, values):
		self._points.extend(points)
		self._values.extend(values)

	@property
	def std(self):
		"""spread of the angles; 0 for fewer than four samples"""
		return np.std(self._values) if len(self._values) > 3 else 0


class Field:
	"""reference dewarp.py:108-141: angle field -> unit direction field."""

	def __init__(self, samples, size, phi0, device=None):
		self._size = size
		if device is None:
			if not torch.cuda.is_available():
				raise _lib.OrigamiB200Error("origami_b200 needs a CUDA device (there is no CPU fallback)")
			device = torch.device("cuda", torch.cuda.current_device())
		self._device = torch.device(device)
		if len(samples.points) <= 2:
			self._interp = None
			f = _lib.Field()
			f.constant, f.phi0 = 1, float(phi0)
			self._struct = f
		else:
			self._interp = lininterp(
				np.asarray(samples.points, dtype=np.float64), np.asarray(samples.values, dtype=np.float64),
				(0, 0, size[0], size[1]), device=self._device)
			self._struct = self._interp.field_struct()

	@property
	def struct(self):
		return self._struct

	def march(self, start, n_steps, step):
		"""start (P, 2) float32 -> (n_steps, P, 2) float32: positions before each of n_steps Euler steps
		`pts += self.get(pts) * step` (one launch)."""
		start = np.ascontiguousarray(start, dtype=np.float32).reshape(-1, 2)
		P = start.shape[0]
		if P == 0 or n_steps <= 0:
			return np.zeros((max(n_steps, 0), P, 2), dtype=np.float32)
		d_start = torch.from_numpy(start).to(self._device)
		out = torch.empty((n_steps, P, 2), dtype=torch.float32, device=self._device)
		ops._call("og_field_march", d_start, ([out], lambda st: (
			_lib.ctypes.byref(self._struct), d_start.data_ptr(), P, int(n_steps), float(step), out.data_ptr(), st)))
		return out.cpu().numpy()

	def get(self, pts):
		"""(n, 2) points -> (n, 2) float32 unit vectors (cos phi, sin phi)."""
		pts = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 2)
		if self._interp is None:
			a = np.full(len(pts), self._struct.phi0)
		else:
			a = self._interp(pts.astype(np.float64))
		return np.squeeze(np.dstack([np.cos(a, dtype=np.float32), np.sin(a, dtype=np.float32)]), axis=0)

	def estimate_extent(self, axis, limit, step_size):
		pts = np.array([[0, y] for y in range(0, self._size[1 - axis], step_size)]).astype(np.float32)
		if axis != 0:
			pts = np.flip(pts, axis=-1)
		max_steps = 2 * (1 + self._size[axis] // 1)
		# the whole trajectory in one launch (in batches), the first step count at which every point has passed `limit`
		n_steps, batch = 1, 256
		cur = np.ascontiguousarray(pts)
		done = 0
		while done < max_steps:
			traj = self.march(cur, batch + 1, step_size)  # traj[k] = positions after `done + k` steps
			for k in range(batch):
				if not np.any(traj[k][:, axis] < limit):
					return n_steps
				n_steps += 1
				if n_steps >= max_steps:
					raise RuntimeError("n_steps exceeded %d" % max_steps)
			cur = traj[batch]
			done += batch
		raise RuntimeError("n_steps exceeded %d" % max_steps)


def extrapolate(a, b, x):
	"""the point x beyond b on the ray from a through b"""
	d = np.subtract(b, a)
	unit = d / np.linalg.norm(d)
	return b + x * unit


class GridFactory:
	def __init__(self, page, samples_h, samples_v, grid_res=25, max_grid_size=1000, num_threads=2, device=None):
		size = page.warped.size if hasattr(page, "warped") else tuple(page)
		self._width = size[0]
		self._height = size[1]
		self._grid_res = grid_res
		self._max_grid_size = max_grid_size
		self._num_threads = num_threads
		self._samples_h = samples_h
		self._samples_v = samples_v
		self._device = device

	@property
	def res(self):
		return self._grid_res

	@property
	def std(self):
		return max(self._samples_h.std, self._samples_v.std)

	@cached_property
	def field_h(self):
		return Field(self._samples_h, (self._width, self._height), phi0=0, device=self._device)

	@cached_property
	def field_v(self):
		return Field(self._samples_v, (self._width, self._height), phi0=math.pi / 2, device=self._device)

	@cached_property
	def grid_shape(self):
		est_width = self.field_h.estimate_extent(0, self._width, step_size=self._grid_res)
		est_height = self.field_v.estimate_extent(1, self._height, step_size=self._grid_res)
		if max(est_width, est_height) > self._max_grid_size:
			raise RuntimeError("estimated grid too big: (%d, %d)" % (est_height, est_width))
		return est_height, est_width

	def _extend(self, grid_hv, field, start, cond, step):
		"""Border extension (reference :341-419): march the border points outwards until all have left
		the page; returns the list of new borders (without the starting one)."""
		max_borders = self._max_grid_size // 2
		borders = []
		cur = np.ascontiguousarray(start, dtype=np.float32)
		if not cond(cur):
			return borders
		batch = 16
		while True:
			traj = field.march(cur, batch + 1, step)
			for k in range(1, batch + 1):
				if len(borders) + 1 >= max_borders:
					raise RuntimeError("border extension not terminating")
				borders.append(traj[k])
				if not cond(traj[k]):
					return borders
			cur = traj[batch]

	def extend_border_h(self, grid_hv, side):
		if side == "left":
			new = self._extend(grid_hv, self.field_h, grid_hv[:, 0], lambda x: np.any(x[:, 0] > 0), -self._grid_res)
			cols = [b.reshape(grid_hv.shape[0], 1, 2) for b in reversed(new)] + [grid_hv]
		elif side == "right":
			new = self._extend(grid_hv, self.field_h, grid_hv[:, -1], lambda x: np.any(x[:, 0] < self._width), self._grid_res)
			cols = [grid_hv] + [b.reshape(grid_hv.shape[0], 1, 2) for b in new]
		else:
			raise ValueError(side)
		return np.hstack(cols) if new else grid_hv

	def extend_border_v(self, grid_hv, side):
		if side == "top":
			new = self._extend(grid_hv, self.field_v, grid_hv[0, :], lambda x: np.any(x[:, 1] > 0), -self._grid_res)
			rows = [b.reshape(1, grid_hv.shape[1], 2) for b in reversed(new)] + [grid_hv]
		elif side == "bottom":
			new = self._extend(grid_hv, self.field_v, grid_hv[-1, :], lambda x: np.any(x[:, 1] < self._height), self._grid_res)
			rows = [grid_hv] + [b.reshape(1, grid_hv.shape[1], 2) for b in new]
		else:
			raise ValueError(side)
		return np.vstack(rows) if new else grid_hv

	@cached_property
	def grid_h(self):
		gh, gw = self.grid_shape
		start = np.array([[0, gy * self._grid_res] for gy in range(gh)], dtype=np.float32)
		traj = self.field_h.march(start, gw, self._grid_res)  # (gw, gh, 2)
		grid = np.ascontiguousarray(np.transpose(traj, (1, 0, 2)))
		assert not np.any(np.isnan(grid))
		return grid

	@cached_property
	def grid_hv(self):
		grid_h = self.grid_h.copy()
		gh, gw = grid_h.shape[:2]
		grid_res = self._grid_res
		# BatchIntersections.make_row_slice pushes the two end points of every h-line below the first
		# outwards (in place in the reference, dewarp.py:198-216), so that rays always find a crossing
		large = gw * grid_res
		for gy in range(1, gh):
			row = grid_h[gy]
			row[0] = extrapolate(row[1].copy(), row[0].copy(), large)
			row[-1] = extrapolate(row[-2].copy(), row[-1].copy(), large)
		dev = self.field_v._device
		d_grid_h = torch.from_numpy(grid_h).to(dev)
		d_out = torch.zeros((gh, gw, 2), dtype=torch.float32, device=dev)
		status = torch.zeros((1,), dtype=torch.int32, device=dev)
		slices = [slice(0, gw)] if self._num_threads < 2 else list(make_slices(gw, self._num_threads))
		for sel in slices:
			max_dy = 0
			for gy in range(1, gh):
				max_dy = max(max_dy, np.max(grid_h[gy, sel, 1]) - np.min(grid_h[gy - 1, sel, 1]))
			max_step_len = max_dy / math.cos(60 * (math.pi / 180))
			c0, c1, _ = sel.indices(gw)
			ops._call("og_grid_vlines", d_grid_h, ([d_out, status], lambda st: (
				_lib.ctypes.byref(self.field_v.struct), d_grid_h.data_ptr(), gh, gw, c0, c1 - c0, float(max_step_len),
				d_out.data_ptr(), status.data_ptr(), st)))
		failed = int(status.item())
		if failed:
			raise RuntimeError("failed to find intersection for i=%d" % (failed - 1))
		grid_hv = d_out.cpu().numpy()
		grid_hv = self.extend_border_h(grid_hv, "left")
		grid_hv = self.extend_border_h(grid_hv, "right")
		grid_hv = self.extend_border_v(grid_hv, "top")
		grid_hv = self.extend_border_v(grid_hv, "bottom")
		assert not np.any(np.isnan(grid_hv))
		return grid_hv


def make_slices(n, k):
	"""range(n) cut into k runs of equal length (the last one shorter), as slices"""
	run = -(-n // k)
	return (slice(lo, lo + run) for lo in range(0, n, run))
