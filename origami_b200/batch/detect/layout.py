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


def regions_binarized(page, median_line_height, separator_labels=(), on_device=False):
	"""page: origami_b200.core.page.Page built with a Grid; separator_labels: iterable of
	(labels uint8 ndarray, background label value) for every SEPARATOR prediction.
	Returns the float32 (Hd, Wd) array of Regions.binarized; on_device=True leaves it on the GPU as a
	CUDA tensor (what SplitDetector.detect_regions takes, so the page-sized image never travels)."""
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
	out = ops.or_dilate3(binary, mask_total)
	return out if on_device else ops.to_host(out)


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

	def _peaks(self, column_quantiles, scale):
		import scipy.ndimage
		import scipy.signal
		ink_h = scipy.ndimage.convolve(column_quantiles, kernel(int(self._smooth * scale)), mode="nearest")
		span = int(self._border * len(ink_h))
		ink_h[:span] = 0
		ink_h[-span:] = 0
		return scipy.signal.find_peaks(-ink_h, height=-self._intensity, distance=int(self._width * scale))

	def detect_regions(self, binarized, bounds, scales, transposed=False):
		"""The detector on the crops of all contours of a page at once (what FixSpillOverH / FixSpillOverV,
		layout.py:947-1075, do one contour at a time): binarized is Regions.binarized as a CUDA tensor (or a
		NumPy array, uploaded once), bounds a list of contour bounds (minx, miny, maxx, maxy), scales the
		line height per contour; transposed=True is FixSpillOverV's crop.transpose().  One kernel launch
		for all periodograms; returns [(peaks, info, (minx, miny))] like crop() + __call__."""
		import torch
		if not isinstance(binarized, torch.Tensor):
			binarized = torch.from_numpy(np.ascontiguousarray(binarized)).to("cuda")
		H, W = int(binarized.shape[0]), int(binarized.shape[1])
		boxes = []
		for minx, miny, maxx, maxy in bounds:  # crop(): layout.py:886-894
			y0, x0 = int(max(0, miny)), int(max(0, minx))
			y1, x1 = int(min(maxy, H)), int(min(maxx, W))
			boxes.append((x0, y0, max(x1, x0), max(y1, y0)))
		cols = ops.region_psd_quantile(binarized, boxes, transposed=transposed, quantile=self._quantile)
		scales = np.broadcast_to(np.asarray(scales, dtype=np.float64), (len(boxes),))
		res = []
		for box, cq, scale in zip(boxes, cols, scales):
			peaks, info = self._peaks(np.array(cq), scale) if len(cq) else (np.zeros(0, np.int64), {"peak_heights": np.zeros(0)})
			res.append((peaks, info, (box[0], box[1])))
		return res

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
