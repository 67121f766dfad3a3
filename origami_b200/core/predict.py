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
