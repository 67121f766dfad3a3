"""
ORACLE (test infrastructure).  Imports the UNMODIFIED reference modules from
/root/reference behind a sys.modules shim for the third-party packages that are not
installed in this image (shapely, scikit-image, cairocffi, cached_property).  Only
usable in the build container: /root/reference does not exist on the GPU box, so
nothing that runs there (-m gpu tests, smoke(), bench.py) may import this module.
It is used by oracle/make_golden.py to produce tests/golden/*.npz and by the
not-gpu tests that re-check those vectors when the reference tree is present.
"""

import functools
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"


def available():
	return os.path.isdir(os.path.join(REFERENCE_ROOT, "origami", "core"))


def _stub(name, **attrs):
	m = types.ModuleType(name)
	m.__dict__.update(attrs)
	sys.modules[name] = m
	return m


def install_shim():
	if "cached_property" not in sys.modules:
		_stub("cached_property", cached_property=functools.cached_property)
	if "shapely" not in sys.modules:
		sh = _stub("shapely")
		for sub in ("geometry", "ops", "wkt", "affinity"):
			setattr(sh, sub, _stub("shapely." + sub))
		# origami/core/block.py:466 builds a TextAreaFactory() at import time
		sh.strtree = _stub("shapely.strtree", STRtree=lambda geoms: None)
	if "skimage" not in sys.modules:
		sk = _stub("skimage")
		sk.filters = _stub("skimage.filters")
		sk.morphology = _stub("skimage.morphology")
	if "cairocffi" not in sys.modules:
		_stub("cairocffi")
	if REFERENCE_ROOT not in sys.path:
		sys.path.insert(0, REFERENCE_ROOT)


def load():
	"""Returns (Grid, Dewarper, Line) from the reference."""
	install_shim()
	from origami.core.dewarp import Grid, Dewarper
	from origami.core.block import Line
	return Grid, Dewarper, Line


class _Stage:
	is_dewarped = True


class _Page:
	def __init__(self, warped_pil, dewarper):
		self.warped = warped_pil
		self.dewarper = dewarper


class _Block:
	def __init__(self, page):
		self.page = page
		self.stage = _Stage()


def make_line(Line, block, p, right, up):
	"""A reference Line without its shapely-dependent constructor
	(origami/core/block.py:36-53): the fields the hot path reads are set directly."""
	import numpy as np
	ln = Line.__new__(Line)
	ln._p = np.array(p)
	ln._right = np.array(right)
	ln._up = np.array(up)
	ln._block = block
	ln._confidence = 1
	return ln
