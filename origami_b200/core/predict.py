"""
Label argmax of the segmentation post-process (reference origami/core/predict.py:281-285):
labels = argmax_c sum_m probs[m, y, x, c] as uint8, first maximum wins; and the per-class
masks of origami/core/contours.py:483-489.
"""

import numpy as np
import torch

from .. import ops


def argmax_labels(pr_masks, mask_values=None, device=None):
	"""pr_masks: (M, h, w, C) float32 ndarray / tensor (the list of model.predict outputs,
	stacked) -> labels (h, w) uint8 ndarray [, masks (n, h, w) bool ndarray]."""
	if isinstance(pr_masks, (list, tuple)):
		pr_masks = np.stack([np.asarray(p, dtype=np.float32).squeeze() for p in pr_masks])
	if isinstance(pr_masks, np.ndarray):
		t = torch.from_numpy(np.ascontiguousarray(pr_masks, dtype=np.float32)).to(device or "cuda")
	else:
		t = pr_masks.contiguous()
	if t.dim() == 3:
		t = t.unsqueeze(0)
	labels, masks = ops.argmax_labels(t, mask_values)
	if mask_values is None:
		return labels.cpu().numpy()
	return labels.cpu().numpy(), masks.cpu().numpy().astype(bool)


def class_masks(labels, classes):
	"""dict(c -> (pixels == c.value)) for an iterable of enum-like classes (contours.py:488)."""
	classes = list(classes)
	values = [int(getattr(c, "value", c)) for c in classes]
	pixels = np.ascontiguousarray(labels, dtype=np.uint8)
	masks = ops.class_masks(ops.to_device_image(pixels), values).cpu().numpy().astype(bool)
	return dict((c, masks[i]) for i, c in enumerate(classes))


class Tile:
	"""reference predict.py:43-63."""

	def __init__(self, outer, inner):
		self._outer = outer
		self._inner = inner

	@property
	def outer(self):
		return self._outer

	@property
	def inner(self):
		return self._inner

	def read_outer(self, pixels):
		x0, y0, x1, y1 = self._outer
		return pixels[y0:y1, x0:x1]

	def write_inner(self, labels, data):
		x0, y0, x1, y1 = self._inner
		dx, dy = np.array(self._inner[:2]) - np.array(self._outer[:2])
		labels[y0:y1, x0:x1] = data[dy:dy + (y1 - y0), dx:dx + (x1 - x0)]

	def write_inner_argmax(self, device_labels, pr_masks):
		"""argmax of this tile's summed model outputs (predict.py:281-284) written straight into the
		page's device label map: pr_masks = list of (1, th, tw, C) outputs, or one (M, th, tw, C) tensor."""
		if isinstance(pr_masks, (list, tuple)):
			pr_masks = np.stack([np.asarray(p, dtype=np.float32).reshape(p.shape[-3:]) for p in pr_masks])
		t = pr_masks if isinstance(pr_masks, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(pr_masks, dtype=np.float32)).to(device_labels.device)
		return ops.argmax_tile(t.contiguous(), self._inner, self._outer[:2], device_labels)


class Tiles:
	"""reference predict.py:66-113: overlapping tiles with non-overlapping inner rectangles."""

	def __init__(self, tile_size, beta0=50):
		self._tile_size = tile_size
		assert all(beta0 < s for s in tile_size)
		self._beta0 = beta0

	def _tiles_1(self, full_size, tile_size):
		import math
		if tile_size == full_size:
			yield (0, full_size), (0, full_size)
			return
		n_steps = math.ceil(full_size / tile_size)
		while True:
			r = (full_size - tile_size) / ((n_steps - 1) * tile_size)
			if tile_size * (1 - r) > self._beta0:
				break
			n_steps += 1
		x0, x1 = [], []
		for i in range(n_steps):
			x = round(i * tile_size * r)
			x -= max(0, x + tile_size - full_size)
			x0.append(x)
			x1.append(x + tile_size)
		for i in range(n_steps):
			x0_inner = (x1[i - 1] + x0[i]) // 2 if i > 0 else 0
			x1_inner = (x1[i] + x0[i + 1]) // 2 if i < n_steps - 1 else full_size
			yield (x0[i], x1[i]), (x0_inner, x1_inner)

	def __call__(self, full_size):
		import itertools
		p = [list(self._tiles_1(f, t)) for f, t in zip(full_size, self._tile_size)]
		for x, y in itertools.product(*p):
			(x0, x1), (xi0, xi1) = x
			(y0, y1), (yi0, yi1) = y
			yield Tile((x0, y0, x1, y1), (xi0, yi0, xi1, yi1))
