"""
ORACLE (test infrastructure, not product code) -- the reference's CPU path for the page
geometry, restated call for call on the SAME third-party libraries the reference uses
(OpenCV, SciPy, NumPy), so it can run where /root/reference does not exist (the GPU box).

This is the `cpu_baseline` ("kind": "port") of bench.py and the arm timed by
`bench.py --impl reference`: it keeps the reference's data flow, including its costs --
Grid.points("full") materialises the dense map, every line re-reads the page from PIL
(origami/core/block.py:255), line images are mapped over a ThreadPool(8)
(origami/batch/core/lines.py:148).  scikit-image is not installed, so the two threshold
functions come from oracle/exact.py (NumPy restatement of skimage 0.16.2).

Only tests/, __graft_entry__.smoke() and bench.py may import this module.
"""

import multiprocessing.pool

import cv2
import numpy as np
import PIL.Image
import scipy.interpolate

from . import exact


class RefGrid:
	"""origami/core/dewarp.py:496-576."""

	def __init__(self, hv, res):
		self._hv = np.asarray(hv, dtype=np.float32)
		self._res = int(res)
		self._full = None
		self._inv = None

	def points_full(self):
		# dewarp.py:518-524
		if self._full is None:
			g, s = self._hv, self._res
			h, w = g.shape[:2]
			xs = cv2.resize(g[:, :, 0], (w * s, h * s), interpolation=cv2.INTER_LINEAR)
			ys = cv2.resize(g[:, :, 1], (w * s, h * s), interpolation=cv2.INTER_LINEAR)
			self._full = np.dstack([xs, ys])
		return self._full

	def inverse_yx(self):
		# dewarp.py:561-567
		if self._inv is None:
			full = self.points_full()
			self._inv = scipy.interpolate.RegularGridInterpolator(
				(np.arange(full.shape[0]), np.arange(full.shape[1])), full,
				method="linear", bounds_error=False, fill_value=None)
		return self._inv


def dewarped(page_pil, grid):
	"""Dewarper.dewarped, dewarp.py:608-614."""
	full = grid.points_full()
	return PIL.Image.fromarray(cv2.remap(
		np.array(page_pil), full.astype(np.float32), None, interpolation=cv2.INTER_LINEAR))


def dewarp_image(page_pil, grid, im_pil, interpolation=cv2.INTER_LINEAR):
	"""Dewarper.dewarp_image, dewarp.py:616-627, on a private copy of the map."""
	w0, h0 = page_pil.size
	w1, h1 = im_pil.size
	full = grid.points_full().copy()
	full[:, :, 0] *= w1 / w0
	full[:, :, 1] *= h1 / h0
	return PIL.Image.fromarray(cv2.remap(np.array(im_pil), full.astype(np.float32), None, interpolation=interpolation))


def warped_grid(grid, p, right, up, xsteps=None, ysteps=None, xres=1, yres=1, column=None):
	"""Line.warped_grid, block.py:223-248."""
	p0, right_c, up_c, xres = exact.line_position(p, right, up, xres, column)
	if xsteps is None or ysteps is None:
		rough = warped_grid(grid, p, right, up, xsteps=2, ysteps=2)
	if xsteps is None:
		xsteps = np.max(np.abs(rough[:, 0, 0] - rough[:, 1, 0]))
		xsteps = max(2, int(np.ceil(xsteps * xres)))
	if ysteps is None:
		ysteps = np.max(np.abs(rough[0, :, 1] - rough[1, :, 1]))
		ysteps = max(2, int(np.ceil(ysteps * yres)))
	ys = np.linspace([0, 0], up_c, ysteps)
	xs = np.linspace([0, 0], right_c, xsteps)
	dg = (ys + p0)[:, np.newaxis] + xs[np.newaxis, :]
	dg = np.flip(dg, axis=-1)
	wg = grid.inverse_yx()(dg.reshape((len(ys) * len(xs), 2)))
	return wg.reshape((len(ys), len(xs), 2)).astype(np.float32)


def line_image(page_pil, grid, p, right, up, target_height=48, column=None, binarize=None):
	"""Line.image(dewarped=True), block.py:129-143 + 250-259."""
	wg = warped_grid(grid, p, right, up, ysteps=target_height, column=column)
	pixels = np.array(page_pil)  # block.py:255 copies the page for every line
	pixels = cv2.remap(pixels, wg, None, cv2.INTER_LINEAR)
	pixels = pixels[::-1, :]
	if binarize is not None:
		pixels = exact.binarize(pixels, binarize)
	return pixels


def extract_lines(page_pil, grid, lines, target_height=48, binarize=None, threads=8):
	"""LineExtractor.__call__, batch/core/lines.py:131-149: ThreadPool(8).map over line parts.
	lines: [dict(p, right, up, column)]."""
	def one(ln):
		return line_image(page_pil, grid, ln["p"], ln["right"], ln["up"], target_height, ln.get("column"), binarize)

	# The dense map is built here, in the calling thread, not by whichever pool thread asks first:
	# cv2.setUseOptimized() is per thread for the IPP dispatch, so a map built inside a worker would
	# come from OpenCV's IPP path even when the caller pinned the generic one (they differ by <= 2 ulp).
	grid.inverse_yx()
	with multiprocessing.pool.ThreadPool(processes=threads) as pool:
		return pool.map(one, lines)


def confidence_counts(grid, labels, page_shape, p, right, up, n_classes, res=0.5):
	"""ConfidenceSampler.__call__, batch/detect/lines.py:34-54."""
	g = warped_grid(grid, p, right, up, xres=res, yres=res)
	h0, w0 = page_shape
	h1, w1 = labels.shape
	g[:, :, 0] *= w1 / w0
	g[:, :, 1] *= h1 / h0
	lab = cv2.remap(labels, g, None, cv2.INTER_NEAREST)
	return np.bincount(lab.flatten(), minlength=n_classes)


def argmax_labels(pr_masks):
	"""predict.py:281-284."""
	pr_mask = np.sum(pr_masks, axis=0)
	return np.argmax(pr_mask.squeeze(), axis=-1).astype(np.uint8)


def page_path(page, grid_hv, cell, lines, target_height=48, binarize=None, max_lines=None):
	"""One page through the reference path; returns (dewarped ndarray, [strips])."""
	g = RefGrid(grid_hv, cell)
	pil = PIL.Image.fromarray(page)
	dw = np.array(dewarped(pil, g))
	sel = lines if max_lines is None else lines[:max_lines]
	strips = extract_lines(pil, g, sel, target_height, binarize)
	return dw, strips
