"""
Grid and Dewarper with the interface of the reference's origami/core/dewarp.py:496-627,
backed by the sm_100a kernels.  Same names, argument order, defaults and return types
(NumPy arrays / PIL images at the API edge); the sparse grid lives on the GPU and nothing
is computed on the CPU.

Differences that are deliberate:
  * Dewarper.dewarp_image does not scale the cached dense map in place (the reference does,
    origami/core/dewarp.py:621-623, which corrupts later Grid.points("full") users); the
    scale is folded into the kernel instead.
"""

import io
import json
import zipfile
from functools import cached_property

import numpy as np
import PIL.Image
import torch

from .. import ops
from .math import Geometry

INTER_NEAREST = ops.OG_INTER_NEAREST  # == cv2.INTER_NEAREST
INTER_LINEAR = ops.OG_INTER_LINEAR  # == cv2.INTER_LINEAR


def _device(device=None):
	if device is not None:
		return torch.device(device)
	if not torch.cuda.is_available():
		raise ops._lib.OrigamiB200Error("origami_b200 needs a CUDA device (there is no CPU fallback)")
	return torch.device("cuda", torch.cuda.current_device())


class Grid:
	"""Sparse dewarping grid: (gh, gw, 2) fp32 lattice on the dewarped page holding warped-page
	(x, y) positions (reference docs/formats.md:298-306)."""

	def __init__(self, hv, res, device=None):
		hv = np.ascontiguousarray(hv, dtype=np.float32)
		if hv.ndim != 3 or hv.shape[2] != 2:
			raise ValueError("grid must have shape (gh, gw, 2)")
		self._grid_hv = hv
		self._grid_res = int(res)
		self._device_arg = device
		self._points_cache = dict()
		self._device_points = None

	@cached_property
	def _device(self):
		# resolved on first GPU use, so artifacts can be opened / saved on a host without CUDA
		return _device(self._device_arg)

	# -- container (reference :501-511, :528-553)
	@property
	def geometry(self):
		h, w = self._grid_hv.shape[:2]
		r = self._grid_res
		return Geometry(w * r, h * r)

	@property
	def resolution(self):
		return self._grid_res

	@property
	def shape(self):
		return self._grid_hv.shape

	@cached_property
	def device_grid(self):
		return torch.from_numpy(self._grid_hv).to(self._device)

	@cached_property
	def warping(self):
		pts = self.points("sample")
		dy = (pts[1:, :, 0] - pts[:-1, :, 1]).flatten()
		dx = (pts[:, 1:, 1] - pts[:, :-1, 0]).flatten()
		return max(np.std(dx), np.std(dy))

	@staticmethod
	def create(*args, **kwargs):
		"""reference :532-535: the grid of a page from its flow samples (GridFactory)."""
		from .gridfactory import GridFactory
		factory = GridFactory(*args, **kwargs)
		return Grid(factory.grid_hv, factory.res, device=kwargs.get("device"))

	@staticmethod
	def open(path, device=None):
		with zipfile.ZipFile(path, "r") as zf:
			info = json.loads(zf.read("meta.json").decode("utf8"))
			data = io.BytesIO(zf.read("data.npy"))
			grid = np.load(data, allow_pickle=False)
		grid = grid.reshape(info["shape"])
		return Grid(grid, info["cell"], device=device)

	def save(self, path, compression=zipfile.ZIP_DEFLATED):
		data = io.BytesIO()
		np.save(data, self._grid_hv.astype(np.float32), allow_pickle=False)
		info = dict(version=1, cell=self._grid_res, shape=self._grid_hv.shape)
		with zipfile.ZipFile(path, "w", compression) as zf:
			zf.writestr("data.npy", data.getvalue())
			zf.writestr("meta.json", json.dumps(info))

	# -- A2
	def device_points(self):
		"""Dense map as a CUDA tensor (Hd, Wd, 2) fp32.  Kept on the device once computed -- the
		reference keeps its host copy the same way (lru_cache on Grid.points, dewarp.py:514) -- because
		every line of the page interpolates in it; release_device_points() drops it."""
		if self._device_points is None:
			self._device_points = ops.expand_map(self.device_grid, self._grid_res)
		return self._device_points

	def release_device_points(self):
		self._device_points = None

	def points(self, resolution="full", interpolation=INTER_LINEAR):
		if resolution == "sample":
			return self._grid_hv
		elif resolution == "full":
			if interpolation != INTER_LINEAR:
				raise ValueError("only INTER_LINEAR map expansion is implemented")
			key = (resolution, interpolation)
			if key not in self._points_cache:  # the reference lru_caches 8 of these
				self._points_cache[key] = ops.to_host(self.device_points())
			return self._points_cache[key]
		else:
			raise ValueError(resolution)

	@cached_property
	def transformer(self):
		"""reference :555-559: Transformer over Grid.points("full")[::r, ::r].  Those nodes are read off
		the device (the A5 lookup at an integral pixel returns the dense-map node itself), so the
		281 MB map is not materialised for them."""
		r = self._grid_res
		gh, gw = self._grid_hv.shape[:2]
		ii, jj = np.mgrid[0:gh, 0:gw]
		pts_yx = np.stack([ii.ravel() * r, jj.ravel() * r], axis=1).astype(np.float64)
		nodes = self.inverse_yx(pts_yx).astype(np.float32).reshape(gh, gw, 2)
		return Transformer(nodes, r, device=self._device)

	# -- A5
	@cached_property
	def inverse_yx(self):
		def f(pts_yx):
			pts = np.asarray(pts_yx, dtype=np.float64)
			flat = np.ascontiguousarray(pts.reshape(-1, 2))
			out = ops.inverse_yx(self.device_grid, self._grid_res, torch.from_numpy(flat).to(self._device))
			return out.cpu().numpy().reshape(pts.shape[:-1] + (2,))

		return f

	@cached_property
	def inverse(self):
		interp = self.inverse_yx

		def f(pts):
			return interp(np.flip(pts, axis=-1))

		return f


class Transformer:
	"""Warped -> dewarped point transform (reference origami/core/dewarp.py:143-161): linear
	interpolation of the dewarped lattice positions over the warped positions of the lattice nodes,
	with lingrid's extrapolation outside their hull.  grid: (h, w, 2) warped (x, y) of the node at
	dewarped (j * grid_res, i * grid_res)."""

	def __init__(self, grid, grid_res, device=None):
		from .lingrid import lininterp
		grid = np.asarray(grid)
		h, w = grid.shape[:2]
		source = grid.reshape((h * w, 2))
		target = np.flip(np.dstack(
			np.mgrid[0:h, 0:w]), axis=-1).reshape((h * w, 2)) * grid_res
		minx = np.min(source[:, 0])
		miny = np.min(source[:, 1])
		maxx = np.max(source[:, 0])
		maxy = np.max(source[:, 1])
		self._interp = lininterp(source, target, (minx, miny, maxx, maxy), device=device)

	def transform_points(self, pts):
		"""(n, 2) warped (x, y) -> (n, 2) dewarped (x, y), one kernel launch for all points."""
		pts = np.asarray(pts, dtype=np.float64).reshape(-1, 2)
		if pts.shape[0] == 0:
			return pts.copy()
		out = self._interp(pts)
		assert not np.any(np.isnan(out))
		return out

	def __call__(self, x, y):
		pts = self.transform_points(np.squeeze(np.dstack([x, y]), axis=0))
		return pts[:, 0], pts[:, 1]


class Dewarper:
	def __init__(self, im, grid, pixels=None):
		self._im = im
		self._grid = grid
		self._pixels = pixels  # optional ndarray view of `im` (saves a PIL -> bytes round trip)
		self._pending = None

	@property
	def grid(self):
		return self._grid

	@cached_property
	def device_image(self):
		"""The warped page as a pitched CUDA uint8 image (uploaded once)."""
		pixels = self._pixels if self._pixels is not None else np.asarray(self._im, dtype=np.uint8)
		return ops.to_device_image(pixels, self._grid._device)

	def device_dewarped(self, map_out=None):
		"""The dewarped page on the device.  The fused kernel also emits the grid's dense map when the
		grid does not hold it yet (one launch for both), so that the line kernels of the same page can
		gather from it."""
		grid = self._grid
		if map_out is None and grid._device_points is None:
			gh, gw = grid.shape[:2]
			map_out = torch.empty((gh * grid.resolution, gw * grid.resolution, 2), dtype=torch.float32, device=grid._device)
			dst, _ = ops.remap_page(self.device_image, grid.device_grid, grid.resolution, map_out=map_out)
			grid._device_points = map_out
			return dst
		dst, _ = ops.remap_page(self.device_image, grid.device_grid, grid.resolution, map_out=map_out)
		return dst

	def start_dewarped(self):
		"""Launches the remap and the download of the dewarped page without waiting for either; the
		`dewarped` property collects it.  Page.__init__ calls this, so the copy runs beside whatever
		the caller does next (the line kernels of the same page, typically)."""
		if self._pending is None and "dewarped" not in self.__dict__:
			self._pending = ops.HostCopy(self.device_dewarped())

	@cached_property
	def dewarped(self):
		self.start_dewarped()
		pending, self._pending = self._pending, None
		return PIL.Image.fromarray(pending.result())

	def dewarp_image(self, im, interpolation=INTER_LINEAR):
		w0, h0 = self._im.size
		w1, h1 = im.size
		pixels = np.array(im)
		if pixels.ndim != 2 or pixels.dtype != np.uint8:
			raise ValueError("dewarp_image expects an 8-bit single-channel image")
		src = ops.to_device_image(pixels, self._grid._device)
		dst, _ = ops.remap_page(src, self._grid.device_grid, self._grid.resolution,
			scale=(w1 / w0, h1 / h0), interpolation=interpolation)
		return PIL.Image.fromarray(ops.to_host(dst))
