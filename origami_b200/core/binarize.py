"""
Binarizers with the interface of the reference's origami/core/binarize.py:11-43:
otsu(), sauvola(window_size) and from_string(spec) return callable(PIL.Image) -> PIL.Image
giving (pixels > threshold) * 255.  The thresholds are computed by the CUDA kernels.
"""

import re
from functools import partial

import numpy as np
import PIL.Image

from .. import ops


def _binarize_otsu(im):
	dst, _ = ops.otsu(ops.to_device_image(np.array(im)))
	return PIL.Image.fromarray(dst.cpu().numpy())


def _binarize_sauvola(im, window_size):
	dst = ops.sauvola(ops.to_device_image(np.array(im)), window_size)
	return PIL.Image.fromarray(dst.cpu().numpy())


def otsu():
	f = partial(_binarize_otsu)
	f.spec = ("otsu",)
	return f


def sauvola(window_size=15):
	f = partial(_binarize_sauvola, window_size=window_size)
	f.spec = ("sauvola", int(window_size))
	return f


_SPEC = re.compile(r"^\s*(otsu|sauvola)\s*(?:\(\s*(?:(?:window_size\s*=\s*)?(\d+))?\s*\))?\s*$")


def from_string(spec):
	"""Parses "otsu", "sauvola", "sauvola(window_size=31)", "sauvola(31)" -- the forms the
	reference's build_func_from_string (origami/core/utils.py:13-27) evaluates -- without eval."""
	m = _SPEC.match(spec)
	if not m:
		raise ValueError("unsupported binarizer spec %r" % spec)
	if m.group(1) == "otsu":
		return otsu()
	return sauvola(int(m.group(2))) if m.group(2) else sauvola()
