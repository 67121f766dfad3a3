"""
Page with the interface of the reference's origami/core/page.py:81-117: opens the page as
8-bit gray, eagerly dewarps it when a Grid is given, lazily binarizes with Sauvola-15.
A page can also be built from an in-memory array (the synthetic benchmark pages).
"""

from functools import cached_property
from pathlib import Path

import numpy as np
import PIL.Image

from .binarize import sauvola
from .dewarp import Dewarper
from .math import Geometry


def _find_image_path(path):
	path = Path(path)
	if path.exists():
		return path
	candidates = [c for c in path.parent.glob(path.stem + ".*") if c.is_file() and c.suffix.lower() in (
		".jpg", ".jpeg", ".png", ".tif", ".tiff", ".bmp", ".gif")]
	if len(candidates) != 1:
		raise FileNotFoundError(path)
	return candidates[0]


class Page:
	def __init__(self, path, dewarping_transform=None):
		pixels = None
		if isinstance(path, PIL.Image.Image):
			self._warped = path.convert("L")
		elif isinstance(path, np.ndarray):
			pixels = np.ascontiguousarray(path, dtype=np.uint8)
			self._warped = PIL.Image.fromarray(pixels)  # shares the array's memory
		else:
			self._warped = PIL.Image.open(str(_find_image_path(path))).convert("L")

		if dewarping_transform is not None:
			# the reference dewarps eagerly here (page.py:86-88); so does this, but without waiting: the
			# kernel and the download are queued, `dewarped` collects the image when somebody asks for it
			self._dewarper = Dewarper(self._warped, dewarping_transform, pixels=pixels)
			self._dewarper.start_dewarped()
		else:
			self._dewarper = None

	@property
	def warped(self):
		return self._warped

	@property
	def dewarped(self):
		return self._dewarper.dewarped if self._dewarper is not None else None

	@property
	def _dewarped(self):
		return self.dewarped

	@cached_property
	def binarized(self):
		binarizer = sauvola()
		return binarizer(self.warped)

	def size(self, dewarped):
		return (self._dewarped if dewarped else self._warped).size

	def geometry(self, dewarped):
		return Geometry(*self.size(dewarped))

	def pixels(self, dewarped):
		return np.array(self._dewarped if dewarped else self._warped)

	@property
	def dewarper(self):
		return self._dewarper
