"""CPU tests of the C-ABI library: it loads, exports every declared symbol, and rejects bad
arguments before touching the GPU.  No compute calls are made here."""

import ctypes
import os
import re

import pytest

from origami_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "origami_b200.h")


def _declared():
	text = open(HEADER).read()
	text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
	return sorted(set(re.findall(r"\b(og_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
	assert os.path.exists(_lib.lib_path()), "run python -m origami_b200.build"


def test_every_declared_symbol_is_exported_and_bound():
	lib = _lib.load()
	declared = _declared()
	assert len(declared) >= 15
	for name in declared:
		assert hasattr(lib, name), name
	assert declared == _lib.EXPORTED


def test_version_and_error_strings():
	lib = _lib.load()
	assert b"sm_100a" in lib.og_version()
	assert isinstance(lib.og_last_error(), bytes)


def test_argument_errors_need_no_gpu():
	lib = _lib.load()
	assert lib.og_expand_map(None, 10, 10, 25, None, None) == -1
	assert b"NULL" in lib.og_last_error()
	buf = (ctypes.c_float * 64)()
	addr = ctypes.addressof(buf)
	assert lib.og_expand_map(addr, 1, 4, 25, addr, None) == -1  # gh < 2
	assert lib.og_expand_map(addr, 2000, 4, 25, addr, None) == -2  # 50000 rows >= 32767
	assert lib.og_sauvola(addr, 8, 8, 8, 4, 0.2, 127.5, addr + 64, 8, None) == -4
	assert b"must be odd" in lib.og_last_error()
	assert lib.og_remap_page(addr, 8, 8, 8, addr, 4, 4, 2, None, 0, None, 1.0, 1.0, 1, None) == -1
	with pytest.raises(ValueError):
		_lib.check(-1)
	with pytest.raises(_lib.OrigamiB200Error):
		_lib.check(700)


def test_line_struct_size_matches_header():
	assert _lib.LINE_DTYPE.itemsize == 112
	assert _lib.LINE_DTYPE.fields["xsteps"][1] == 96
