"""
Line and Block with the interface of the reference's origami/core/block.py on the
page-geometry path: Line.warped_grid / dewarped_image / deskewed_image / image /
warped_path (block.py:129-267).  Geometry that the reference delegates to shapely
(polygons, text areas) is carried as opaque data; the sampling runs on the GPU.

`lines_to_records` / `extract_strips` are the batched forms used by LineExtractor: all
lines of a page go through ONE kernel launch instead of a Python thread pool.
"""

import collections.abc
import enum
import math
from functools import cached_property

import numpy as np
import PIL.Image
import torch

from .. import ops
from .dewarp import INTER_LINEAR


class Stage(enum.Enum):
	"""origami/batch/core/io.py:29-38"""
	WARPED = 0
	DEWARPED = 1
	AGGREGATE = 2
	RELIABLE = 3
	ANY = -1

	@property
	def is_dewarped(self):
		return self.value >= Stage.DEWARPED.value


class Block:
	"""origami/core/block.py:373-392 (the part the line path touches)."""

	def __init__(self, page, polygon, stage):
		self._image_space_polygon = polygon
		self._page = page
		self._stage = stage

	@property
	def page(self):
		return self._page

	@property
	def page_pixels(self):
		return self.page.pixels(self._stage.is_dewarped)

	@property
	def stage(self):
		return self._stage

	@property
	def image_space_polygon(self):
		return self._image_space_polygon

	def _polygon_ring(self):
		"""Exterior ring of the block polygon as an (k, 2) array: accepts a ring, a WKT string or an
		object with .exterior.coords / .bounds (a shapely Polygon)."""
		poly = self._image_space_polygon
		if poly is None:
			return None
		if isinstance(poly, str):
			from .wkt import parse_wkt
			return parse_wkt(poly)[1][0]
		if hasattr(poly, "exterior"):
			return np.asarray(poly.exterior.coords, dtype=np.float64)
		return np.asarray(poly, dtype=np.float64).reshape(-1, 2)

	@cached_property
	def background(self):
		"""0.8-quantile of the block's pixels (block.py:415-420).  The reference extracts them with
		Mask.extract_image(..., background=None), and Mask._extract applies the polygon mask only when a
		background value is given (mask.py:69-71): the quantile runs over the polygon's integer bounding
		box clipped to the page -- no rasteriser is involved."""
		pixels = self.page_pixels
		ring = self._polygon_ring()
		if ring is not None:
			x0, y0 = np.floor(ring.min(axis=0)).astype(np.int64)
			x1, y1 = np.ceil(ring.max(axis=0)).astype(np.int64)
			pixels = pixels[max(y0, 0):max(y1, 0), max(x0, 0):max(x1, 0)]
		return np.quantile(pixels, 0.8)

	def image(self, text_area, background=None):
		"""block.py:396-401: cut-out of the page through the mask of `text_area` (a ring)."""
		from .mask import Mask
		mask = Mask(text_area)
		if background is None:
			background = self.background
		return mask.extract_image(self.page_pixels, background=background)


def position(p, right, up, xres, column):
	"""Line._position (block.py:195-221): clip the baseline to a table column [x0, x1].

	The reference intersects shapely LineStrings: the baseline with the vertical segments
	x = x0 and x = x1 (spanning the baseline's y range +- 1); GEOS evaluates a proper
	intersection in double-double arithmetic and returns an endpoint's own coordinates when the
	other segment passes through it.  Here the parameter of the hit and its ordinate are computed
	in exact rational arithmetic and rounded once, which is the correctly rounded intersection
	point and reproduces both cases (host-side scalar work: a handful of fractions per table cell).
	A miss keeps the end point, as intersect_segments' default does."""
	from fractions import Fraction
	p0 = np.asarray(p, dtype=np.float64)
	right = np.asarray(right, dtype=np.float64)
	up = np.asarray(up, dtype=np.float64)
	if column is not None:
		p1 = p0 + right
		px0, px1 = min(p0[0], p1[0]), max(p0[0], p1[0])
		py0, py1 = min(p0[1], p1[1]), max(p0[1], p1[1])
		x0, x1 = column
		if x0 is None:
			x0 = px0
		if x1 is None:
			x1 = px1

		def hit(xv, default):
			xv = float(xv)
			if not (np.isfinite(xv) and np.isfinite(p0).all() and np.isfinite(p1).all()):
				return default
			ax, ay, bx, by = (Fraction(float(v)) for v in (p0[0], p0[1], p1[0], p1[1]))
			dx = bx - ax
			if dx == 0:
				return default  # parallel (or a degenerate baseline): no single intersection point
			t = (Fraction(xv) - ax) / dx
			if t < 0 or t > 1:
				return default
			y = ay + t * (by - ay)
			if not (Fraction(float(py0 - 1)) <= y <= Fraction(float(py1 + 1))):
				return default
			return np.array([xv, float(y)], dtype=np.float64)

		q0 = hit(x0, p0)
		q1 = hit(x1, p1)
		right = q1 - q0
		p0 = q0
		xres = xres * ((x1 - x0) / (px1 - px0))
	return p0, right, up, xres


class Line:
	def __init__(
		self, block, p, right, up,
		tesseract_data=None,
		wkt=None, text_area=None, confidence=1):

		self._tesseract_data = tesseract_data
		self._block = block

		self._p = np.array(p, dtype=np.float64)
		self._right = np.array(right, dtype=np.float64)
		self._up = np.array(up, dtype=np.float64)

		self._wkt = wkt
		self._polygon = None
		self._confidence = confidence

	@property
	def block(self):
		return self._block

	@property
	def confidence(self):
		return self._confidence

	@property
	def info(self):
		td = self._tesseract_data or dict()
		return dict(
			p=self._p.tolist(),
			right=self._right.tolist(),
			up=self._up.tolist(),
			wkt=self._wkt,
			confidence=self._confidence,
			tesseract_data=dict((k, td.get(k)) for k in ("baseline", "descent", "ascent", "height")))

	@cached_property
	def length(self):
		return np.linalg.norm(self._right)

	@cached_property
	def height(self):
		return np.linalg.norm(self._up)

	def _position(self, xres, column):
		return position(self._p, self._right, self._up, xres, column)

	def dewarped_height(self, dewarper):
		"""reference block.py:326-333: length of the line's `up` vector after dewarping (a line of a
		WARPED-stage block)."""
		assert not self._block.stage.is_dewarped
		p0, up = self._p, self._up
		tfm = dewarper.grid.transformer
		p1 = p0 + up
		tx, ty = tfm(*np.array([p0, p1]).transpose())
		q0, q1 = np.array([tx, ty]).transpose()
		return np.linalg.norm(q1 - q0)

	def record(self, xsteps=None, ysteps=None, xres=1, yres=1, column=None):
		"""The og_line descriptor of this line for the given sampling request."""
		p0, right, up, xres = self._position(xres, column)
		return dict(p=self._p, right=self._right, up=self._up, cp=p0, cright=right,
			xres=float(xres), yres=float(yres), xsteps=xsteps, ysteps=ysteps)

	def _grid(self):
		return self.block.page.dewarper.grid

	def warped_grid(self, xsteps=None, ysteps=None, xres=1, yres=1, column=None):
		grid = self._grid()
		grids, _ = warped_grids(grid, [self.record(xsteps, ysteps, xres, yres, column)])
		# grid is [y, x, p] and p is [x, y].
		return grids[0]

	def dewarped_image(self, target_height=48, column=None, interpolation=INTER_LINEAR):
		assert self.block.stage.is_dewarped
		if interpolation != INTER_LINEAR:
			raise ValueError("only INTER_LINEAR strips are implemented")
		dewarper = self.block.page.dewarper
		strips, _ = extract_strips(dewarper, [self.record(ysteps=target_height, column=column)])
		return PIL.Image.fromarray(strips[0])

	@property
	def image_space_polygon(self):
		"""Exterior ring of the line polygon, (k, 2) float64 (the reference keeps a shapely Polygon
		parsed from the line's wkt, block.py:48-55)."""
		if self._polygon is None and self._wkt:
			from .wkt import parse_wkt
			self._polygon = parse_wkt(self._wkt)[1][0]
		return self._polygon

	def masked_image(self, mode="polygon"):
		"""block.py:143-153."""
		if mode not in ("polygon", "bbox"):
			raise ValueError(mode)
		from .mask import Mask
		mask = Mask(self.image_space_polygon)
		bg = self._block.background if mode == "polygon" else None
		image, pos = mask.extract_image(self._block.page_pixels, background=bg)
		return image

	def _extract_deskewed(self, pixels, target_height=48, background=255, interpolation=None):
		"""block.py:155-186: cv2.getAffineTransform + cv2.warpAffine (INTER_AREA and INTER_LINEAR give
		identical bytes for warpAffine), then the cut-out through the line polygon carried along by the
		same affine map; pixels outside the polygon become `background`.  One upload, two kernels."""
		p, right, up = self._p, self._right, self._up
		width = int(math.ceil(np.linalg.norm(right)))
		M = affine_transform(
			np.array([p, p + right, p + up]).astype(np.float32),
			np.array([(0, target_height - 1), (width, target_height - 1), (0, 0)]).astype(np.float32))
		src = ops.to_device_image(np.asarray(pixels, dtype=np.uint8))
		warped = ops.warp_affine(src, M, width, target_height, int(background))
		ring = self.image_space_polygon
		if ring is not None and len(ring) >= 3:
			# shapely.affinity.affine_transform(polygon, to_shapely_matrix(matrix)), then Mask(..., bounds=(0, 0, width, height))
			moved = ring @ np.asarray(M, dtype=np.float64)[:, :2].T + np.asarray(M, dtype=np.float64)[:, 2]
			ops.polygon_mask(np.round(moved).astype(np.int32), width, target_height, image=warped, background=int(background))
		return warped.cpu().numpy()

	def deskewed_image(self, target_height=48, interpolation=None):
		return PIL.Image.fromarray(self._extract_deskewed(
			self._block.page_pixels, target_height, self._block.background, interpolation))

	@property
	def warped_image(self):
		"""Line.image(dewarped=False, deskewed=False) reads `self.warped_image` (block.py:138-139), an
		attribute the reference never defines; the cut-out of the line polygon (masked_image) is what
		the neighbouring code suggests it should be."""
		return self.masked_image("polygon")

	def image(
		self, target_height=48, column=None,
		dewarped=True, deskewed=True, binarizer=None):

		if dewarped:
			im = self.dewarped_image(target_height, column=column)
		elif deskewed:
			im = self.deskewed_image(target_height)
		else:
			im = self.warped_image

		if binarizer:
			im = binarizer(im)

		return im

	def warped_path(self, resolution=1):
		assert self.block.stage.is_dewarped
		warped_grid = self.warped_grid(ysteps=2, xres=resolution)
		height = np.median(np.linalg.norm(warped_grid[1] - warped_grid[0], axis=-1))
		return np.mean(warped_grid, axis=0), abs(height)


def affine_transform(src_tri, dst_tri):
	"""cv2.getAffineTransform (host-side scalars): OpenCV's own routine when it is installed (same
	LU solve, same last bits as the reference), else the 6x6 system solved in float64 on
	fp32-rounded points (equal to 1e-9)."""
	try:
		import cv2
		return cv2.getAffineTransform(np.asarray(src_tri, dtype=np.float32), np.asarray(dst_tri, dtype=np.float32))
	except ImportError:
		pass
	src = np.asarray(src_tri, dtype=np.float32).astype(np.float64)
	dst = np.asarray(dst_tri, dtype=np.float32).astype(np.float64)
	A = np.zeros((6, 6))
	b = np.zeros(6)
	for i in range(3):
		A[i, 0:2] = src[i]
		A[i, 2] = 1
		A[i + 3, 3:5] = src[i]
		A[i + 3, 5] = 1
		b[i] = dst[i, 0]
		b[i + 3] = dst[i, 1]
	return np.linalg.solve(A, b).reshape(2, 3)


# ---------------------------------------------------------------------------------------
# batched forms
# ---------------------------------------------------------------------------------------

class StripList(collections.abc.Sequence):
	"""The strips of one page as a read-only sequence of (ysteps, xsteps) uint8 views into the
	packed result buffer; a view is only made when an item is accessed."""

	def __init__(self, flat, offsets, steps):
		self._flat, self._offs, self._steps = flat, offsets, steps

	def __len__(self):
		return int(self._steps.shape[0])

	def __iter__(self):
		flat = self._flat
		offs = self._offs.tolist()
		for (w, h), a, b in zip(self._steps.tolist(), offs, offs[1:]):
			yield flat[a:b].reshape(h, w)

	def __getitem__(self, i):
		if isinstance(i, slice):
			return [self[k] for k in range(*i.indices(len(self)))]
		if i < 0:
			i += len(self)
		if not 0 <= i < len(self):
			raise IndexError(i)
		return self._flat[self._offs[i]:self._offs[i + 1]].reshape(self._steps[i, 1], self._steps[i, 0])


class PackedStripList(collections.abc.Sequence):
	"""Binarized strips that travelled as one bit per pixel (og_pack_bits): a read-only sequence of
	(ysteps, xsteps) uint8 arrays of 0 / 255 like StripList, each unpacked when it is accessed.
	`bits` is the packed buffer (numpy.packbits order); strip i starts at bit offsets[i]."""

	def __init__(self, bits, offsets, steps):
		self._bits, self._offs, self._steps = bits, offsets, steps

	def __len__(self):
		return int(self._steps.shape[0])

	def _strip(self, a, b, h, w):
		first = a >> 3
		px = np.unpackbits(self._bits[first:(b + 7) >> 3], count=(b - 8 * first))[a - 8 * first:]
		px *= 255
		return px.reshape(h, w)

	def __iter__(self):
		offs = self._offs.tolist()
		for (w, h), a, b in zip(self._steps.tolist(), offs, offs[1:]):
			yield self._strip(a, b, h, w)

	def __getitem__(self, i):
		if isinstance(i, slice):
			return [self[k] for k in range(*i.indices(len(self)))]
		if i < 0:
			i += len(self)
		if not 0 <= i < len(self):
			raise IndexError(i)
		return self._strip(int(self._offs[i]), int(self._offs[i + 1]), int(self._steps[i, 1]), int(self._steps[i, 0]))


def _plan(grid, records):
	la = records if isinstance(records, np.ndarray) else ops.pack_lines(records)
	dev = grid._device
	ld = ops.lines_to_device(la, dev) if len(records) else torch.zeros((8,), dtype=torch.uint8, device=dev)
	plan = ops.line_steps(grid.device_grid, grid.resolution, ld, len(records))
	steps, offs = ops.plan_to_host(plan)  # one download for step counts and offsets
	if len(records) and (steps <= 0).any():
		bad = int(np.argmax((steps <= 0).any(axis=1)))
		raise ValueError("line %d has a degenerate or non-finite extent" % bad)
	return plan, steps, offs


def warped_grids(grid, records):
	"""Line.warped_grid for a list of og_line records -> ([ (ysteps, xsteps, 2) fp32 ], steps)."""
	plan, steps, offs = _plan(grid, records)
	out = torch.empty((max(int(offs[-1]), 1) * 2,), dtype=torch.float32, device=grid._device)
	ops.warped_grids(grid.device_grid, grid.resolution, plan, out, dense_map=grid._device_points)
	flat = ops.to_host(out)
	res = [flat[2 * offs[i]:2 * offs[i + 1]].reshape(steps[i][1], steps[i][0], 2) for i in range(len(records))]
	return res, steps


def extract_strips(dewarper, records, binarize=None, want_bbox=False):
	"""Line.image(dewarped=True) for a list of records in one launch.

	binarize: None | ("otsu",) | ("sauvola", window).  Returns (sequence of (ysteps, xsteps) uint8
	arrays -- views into one result buffer --, bbox | None).  When the page's dense map is on the
	device (Dewarper.dewarped left it there) the samples gather from it.
	"""
	grid = dewarper.grid
	plan, steps, offs = _plan(grid, records)
	dev = grid._device
	total = max(int(offs[-1]), 1)
	gray = torch.empty((total,), dtype=torch.uint8, device=dev)
	bbox = torch.empty((max(len(records), 1), 4), dtype=torch.int32, device=dev) if want_bbox else None
	ops.extract_lines(dewarper.device_image, grid.device_grid, grid.resolution, plan, gray, bbox, binarize=binarize,
		dense_map=grid._device_points)
	bb = ops.to_host(bbox)[:len(records)] if want_bbox else None
	flat = ops.to_host(gray)
	return StripList(flat, offs, steps), bb
