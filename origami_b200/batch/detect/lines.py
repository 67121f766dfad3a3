"""
ConfidenceSampler with the interface of the reference's origami/batch/detect/lines.py:18-54,
plus a batched form that samples every line of a page in one launch.
"""

import numpy as np
import torch

from ... import ops


def scale_grid(s0, s1, grid):
	"""In-place scaling of a (h, w, 2) fp32 warped grid from page size s0 to label size s1
	(origami/batch/detect/lines.py:18-22)."""
	h0, w0 = s0
	h1, w1 = s1
	grid[:, :, 0] *= w1 / w0
	grid[:, :, 1] *= h1 / h0


class ConfidenceSampler:
	def __init__(self, blocks, segmentation):
		self._predictions = dict()
		for p in segmentation.predictions:
			self._predictions[p.name] = p

		self._page = list(blocks.values())[0].page
		self._page_shape = tuple(reversed(self._page.warped.size))
		self._labels_dev = dict()

	def _device_labels(self, name):
		if name not in self._labels_dev:
			grid = self._page.dewarper.grid
			self._labels_dev[name] = ops.to_device_image(
				np.asarray(self._predictions[name].labels, dtype=np.uint8), grid._device)
		return self._labels_dev[name]

	def counts(self, prediction_name, lines, res=0.5):
		"""(n, n_classes) int64 label histograms of all `lines` (one launch)."""
		predictor = self._predictions[prediction_name]
		grid = self._page.dewarper.grid
		records = [line.record(xres=res, yres=res) for line in lines]
		la = ops.pack_lines(records)
		plan = ops.line_steps(grid.device_grid, grid.resolution, ops.lines_to_device(la, grid._device), len(records))
		n_classes = len(predictor.classes)
		counts, overflow = ops.sample_labels(self._device_labels(prediction_name), self._page_shape,
			grid.device_grid, grid.resolution, plan, n_classes)
		counts = counts.cpu().numpy()
		if int(overflow.sum().item()) != 0:
			raise ValueError("label map holds values >= number of classes")
		return counts

	def evidence(self, prediction_name, counts):
		predictor = self._predictions[prediction_name]
		evidence = dict()
		sum_all = np.sum(counts)
		if sum_all > 0:
			for k in predictor.classes:
				key = "%s/%s" % (prediction_name, k.name)
				evidence[key] = counts[k.value] / sum_all
		return evidence

	def __call__(self, path, line, res=0.5):
		prediction_name, predictor_class = path[:2]
		counts = self.counts(prediction_name, [line], res)[0]
		return self.evidence(prediction_name, counts)

	def sample_all(self, items, res=0.5):
		"""items: [(path, line)] -> [evidence dict], batched per predictor."""
		out = [None] * len(items)
		by_name = dict()
		for i, (path, line) in enumerate(items):
			by_name.setdefault(path[0], []).append(i)
		for name, idx in by_name.items():
			counts = self.counts(name, [items[i][1] for i in idx], res)
			for j, i in enumerate(idx):
				out[i] = self.evidence(name, counts[j])
		return out
