"""
Mask with the interface of the reference's origami/core/mask.py:8-74: a binary polygon mask over
the polygon's integer bounding box, and cut-outs of page pixels through it.  The reference fills
the polygon with cairo; here og_polygon_mask computes the pixel coverage on the GPU (every pixel
covered by at least 1/510 belongs to the mask, as cairo's `R > 0` after 8-bit rounding; cairo's
rasteriser itself is absent and not reproduced bit for bit).

`shape` is an (k, 2) exterior ring or a list of rings (a MultiPolygon's exteriors); only exteriors
are filled, as in the reference (mask.py:38-41).
"""

import numpy as np
import PIL.Image

from .. import ops


def _rings(shape):
	if hasattr(shape, "exterior"):  # a shapely Polygon, should one be handed in
		return [np.asarray(shape.exterior.coords, dtype=np.float64)]
	if hasattr(shape, "geoms"):
		return [np.asarray(g.exterior.coords, dtype=np.float64) for g in shape.geoms]
	if isinstance(shape, np.ndarray):
		return [np.asarray(shape, dtype=np.float64).reshape(-1, 2)]
	return [np.asarray(r, dtype=np.float64).reshape(-1, 2) for r in shape]


class Mask:
	def __init__(self, shape, bounds=None, buffer=0, device=None):
		rings = _rings(shape)
		if not rings or any(len(r) < 3 for r in rings):
			raise ValueError("unsupported shape for mask")
		if bounds is None:
			allpts = np.concatenate(rings, axis=0)
			minx, miny = np.floor(allpts.min(axis=0)).astype(np.int32)
			maxx, maxy = np.ceil(allpts.max(axis=0)).astype(np.int32)
			if buffer > 0:
				minx -= buffer
				miny -= buffer
				maxx += buffer
				maxy += buffer
		else:
			minx, miny, maxx, maxy = bounds
		w = int(maxx - minx)
		h = int(maxy - miny)
		mask = np.zeros((max(h, 0), max(w, 0)), dtype=bool)
		if w > 0 and h > 0:
			for ring in rings:
				pts = np.round(ring - np.array([minx, miny])).astype(np.int32)  # mask.py:39-40
				mask |= ops.polygon_mask(pts, w, h, device=device).cpu().numpy() > 0
		self._mask = mask
		self._bbox = (int(minx), int(miny), w, h)

	@property
	def binary(self):
		return self._mask

	@property
	def bounds(self):
		minx, miny, w, h = self._bbox
		return minx, miny, minx + w, miny + h

	def _extract(self, bbox, pixels, background=255):
		"""The part of the box that lies on the page, the pixels outside the polygon replaced by
		`background` (None: left as they are); returns it with its top-left corner on the page."""
		x, y, w, h = bbox
		rows, cols = pixels.shape[:2]
		x0, y0 = max(x, 0), max(y, 0)
		x1, y1 = min(x + w, cols), min(y + h, rows)
		region = np.array(pixels[y0:y1, x0:x1])
		if background is not None:
			inside = self._mask[y0 - y:y1 - y, x0 - x:x1 - x]
			region[~inside] = background
		return region, (x0, y0)

	def extract(self, pixels, background=255):
		return self._extract(self._bbox, pixels, background)

	def cutout(self, pixels, background=255):
		r, _ = self._extract(self._bbox, pixels, background)
		return r

	def extract_image(self, pixels, background=255):
		cutout, pos = self.extract(pixels, background)
		return PIL.Image.fromarray(cutout), pos
