"""
ctypes binding of liborigami_b200.so (include/origami_b200.h).

There is no CPU fallback: if the library is missing or CUDA is unavailable every compute
entry point raises.  The library is built in-tree by `python -m origami_b200.build`.
"""

import ctypes
import os

import numpy as np

from . import build as _build

_LIB = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_int64 = ctypes.c_int64
c_float = ctypes.c_float
c_double = ctypes.c_double

OG_INTER_NEAREST = 0
OG_INTER_LINEAR = 1
OG_BIN_NONE = 0
OG_BIN_OTSU = 1
OG_BIN_SAUVOLA = 2

# struct og_line (include/origami_b200.h), 112 bytes
LINE_DTYPE = np.dtype([
	("p", np.float64, 2), ("right", np.float64, 2), ("up", np.float64, 2),
	("cp", np.float64, 2), ("cright", np.float64, 2),
	("xres", np.float64), ("yres", np.float64),
	("xsteps", np.int32), ("ysteps", np.int32), ("flags", np.int32), ("pad_", np.int32)],
	align=True)
assert LINE_DTYPE.itemsize == 112

class TriTables(ctypes.Structure):
	"""struct og_tri_tables (include/origami_b200.h)"""
	_fields_ = [("cell_start", c_void_p), ("cell_tris", c_void_p), ("transform", c_void_p), ("simplices", c_void_p),
		("values", c_void_p), ("ox", c_double), ("oy", c_double), ("cell", c_double), ("eps", c_double),
		("gx", ctypes.c_int32), ("gy", ctypes.c_int32)]


class Field(ctypes.Structure):
	"""struct og_field (include/origami_b200.h)"""
	_fields_ = [("inter", TriTables), ("extra", TriTables), ("bounds", c_double * 4), ("phi0", c_double),
		("has_inter", ctypes.c_int32), ("constant", ctypes.c_int32)]


# name -> argtypes; every function returns int except the two string getters
_SIGNATURES = {
	"og_expand_map": [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p],
	"og_remap_page": [c_void_p, c_int, c_int, c_int64, c_void_p, c_int, c_int, c_int,
		c_void_p, c_int64, c_void_p, c_float, c_float, c_int, c_void_p],
	"og_inverse_yx": [c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p],
	"og_line_steps": [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p],
	"og_warped_grids": [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
		c_void_p, c_int64, c_void_p],
	"og_extract_lines": [c_void_p, c_int, c_int, c_int64, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
		c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p],
	"og_binarize_strips": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p],
	"og_pack_strips": [c_void_p, c_void_p, c_void_p, c_int, c_int64, c_int, c_int, c_int, c_int, c_int,
		c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
	"og_pack_bits": [c_void_p, c_int64, c_void_p, c_void_p],
	"og_tri_interp": [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_double, c_double, c_double,
		c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_int, c_void_p],
	"og_field_march": [ctypes.POINTER(Field), c_void_p, c_int, c_int, c_float, c_void_p, c_void_p],
	"og_grid_vlines": [ctypes.POINTER(Field), c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p],
	"og_sauvola": [c_void_p, c_int, c_int, c_int64, c_int, c_double, c_double, c_void_p, c_int64, c_void_p],
	"og_otsu": [c_void_p, c_int, c_int, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
	"og_otsu_noalloc": [c_void_p, c_int, c_int, c_int64, c_void_p, c_int64, c_void_p, c_void_p],
	"og_sample_labels": [c_void_p, c_int, c_int, c_int64, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p,
		c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p],
	"og_argmax_labels": [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p],
	"og_argmax_tile": [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
		c_void_p, c_int, c_int, c_int64, c_int, c_int, c_void_p],
	"og_class_masks": [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_int, c_void_p],
	"og_label_mask_ne": [c_void_p, c_int, c_int, c_int64, c_int, c_void_p, c_int64, c_void_p],
	"og_or_dilate3": [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p],
	"og_column_psd_quantile": [c_void_p, c_int, c_int, c_int, c_int64, c_double, c_void_p, c_void_p],
	"og_region_psd_quantile": [c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_int, c_int, c_int, c_double, c_void_p, c_void_p],
	"og_polygon_mask": [c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_int, c_void_p],
	"og_warp_affine": [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_int, c_int, c_int64, c_int, c_void_p],
}

EXPORTED = sorted(list(_SIGNATURES) + ["og_last_error", "og_version"])


class OrigamiB200Error(RuntimeError):
	pass


def lib_path():
	# ORIGAMI_B200_LIB selects another build of the same library (kernel tuning experiments)
	return os.environ.get("ORIGAMI_B200_LIB") or _build.LIB_PATH


def load():
	"""Loads the shared library (once).  Raises if it has not been built."""
	global _LIB
	if _LIB is not None:
		return _LIB
	path = lib_path()
	if not os.path.exists(path):
		raise OrigamiB200Error(
			"%s is missing: run `python -m origami_b200.build` (there is no CPU fallback)" % path)
	lib = ctypes.CDLL(path)
	for name, argtypes in _SIGNATURES.items():
		fn = getattr(lib, name)
		fn.argtypes = argtypes
		fn.restype = c_int
	lib.og_last_error.restype = ctypes.c_char_p
	lib.og_last_error.argtypes = []
	lib.og_version.restype = ctypes.c_char_p
	lib.og_version.argtypes = []
	_LIB = lib
	return lib


def check(rc):
	"""Maps the C return convention onto the exceptions the reference raises: argument
	errors -> ValueError (cv2.error / ValueError in the reference), CUDA errors -> RuntimeError."""
	if rc == 0:
		return
	msg = load().og_last_error().decode("utf8", "replace")
	if rc < 0:
		raise ValueError("origami_b200: %s (code %d)" % (msg, rc))
	raise OrigamiB200Error("origami_b200: CUDA error %d: %s" % (rc, msg))


def call(name, *args):
	check(getattr(load(), name)(*args))
