"""
Pixel helper of the layout stage with the behaviour of the reference's Regions.binarized
(origami/batch/detect/layout.py:157-186): Sauvola on the dewarped page with a window derived
from the median line height, OR-ed with the 3x3-dilated, dewarped separator masks -> float32 0/1.
The rule engine of the layout stage itself is out of scope.
"""

import numpy as np

from ... import ops
from ...core.dewarp import INTER_NEAREST


def sauvola_window(median_line_height):
	"""layout.py:162-166: half the median line height, made odd, at least 3."""
	window_size = int(median_line_height) // 2
	if window_size % 2 == 0:
		window_size += 1
	return max(window_size, 3)


def regions_binarized(page, median_line_height, separator_labels=()):
	"""page: origami_b200.core.page.Page built with a Grid; separator_labels: iterable of
	(labels uint8 ndarray, background label value) for every SEPARATOR prediction.
	Returns the float32 (Hd, Wd) array of Regions.binarized."""
	dewarper = page.dewarper
	gray = dewarper.device_dewarped()
	binary = ops.sauvola(gray, sauvola_window(median_line_height))
	hd, wd = int(gray.shape[0]), int(gray.shape[1])
	w0, h0 = page.warped.size
	grid = dewarper.grid
	mask_total = None
	for labels, background in separator_labels:
		lab = ops.to_device_image(np.ascontiguousarray(labels, dtype=np.uint8), grid._device)
		mask = ops.label_mask_ne(lab, background)
		h1, w1 = int(lab.shape[0]), int(lab.shape[1])
		dmask, _ = ops.remap_page(mask, grid.device_grid, grid.resolution, scale=(w1 / w0, h1 / h0), interpolation=INTER_NEAREST)
		# several separator predictions: OR them before the single dilation (dilation distributes over OR)
		mask_total = dmask if mask_total is None else (mask_total | dmask)
	assert mask_total is None or tuple(mask_total.shape) == (hd, wd)
	return ops.to_host(ops.or_dilate3(binary, mask_total))


def crop(pixels, bounds):
	"""layout.py:886-894 with the contour given by its bounds (minx, miny, maxx, maxy)."""
	minx, miny, maxx, maxy = bounds
	miny = int(max(0, miny))
	minx = int(max(0, minx))
	maxy = int(min(maxy, pixels.shape[0]))
	maxx = int(min(maxx, pixels.shape[1]))
	return pixels[miny:maxy, minx:maxx], (minx, miny)


def kernel(*s):
	"""layout.py:58-59"""
	return np.ones(s) / np.prod(s)


class SplitDetector:
	"""layout.py:915-944: finds the x position of a vertical whitespace gap inside a text region from
	the column-wise power spectra of its (binarized) pixels.  The periodogram and the quantile over
	frequencies run on the GPU for all columns of the crop at once; the smoothing and the peak search
	work on one number per column and stay on the host with the reference's SciPy calls."""

	def __init__(self, quantile=0.9, smooth=1, intensity=0.05, width=2, border=0.1):
		self._quantile = quantile
		self._smooth = smooth
		self._intensity = intensity
		self._width = width
		self._border = border

	def column_quantiles(self, pixels, device=None):
		import torch
		if isinstance(pixels, torch.Tensor):
			t = pixels
		else:
			if pixels.dtype != np.uint8:
				assert pixels.dtype == np.float32
			t = torch.from_numpy(np.ascontiguousarray(pixels)).to(device or "cuda")
		return ops.column_psd_quantile(t, self._quantile).cpu().numpy()

	def __call__(self, pixels, scale, device=None):
		import scipy.ndimage
		import scipy.signal
		ink_h = scipy.ndimage.convolve(
			self.column_quantiles(pixels, device),
			kernel(int(self._smooth * scale)),
			mode="nearest")
		span = int(self._border * len(ink_h))
		ink_h[:span] = 0
		ink_h[-span:] = 0
		peaks, info = scipy.signal.find_peaks(
			-ink_h, height=-self._intensity, distance=int(self._width * scale))
		return peaks, info
