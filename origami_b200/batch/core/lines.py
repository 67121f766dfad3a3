"""
LineRewriter and LineExtractor with the interface of the reference's
origami/batch/core/lines.py:51-149.  The reference maps Line.image over a ThreadPool(8);
here every line of the page goes through one batched kernel launch.
"""

import collections.abc
import logging

import PIL.Image

from ... import ops
from ...core import binarize as _binarize
from ...core.block import extract_strips


class LineRewriter:
	def __init__(self, tables):
		self._columns = dict(
			(tuple(k.split("/")), xs)
			for k, xs in tables["columns"].items())

	def _column_path(self, path, column):
		assert column >= 1
		predictor, label = path[:2]
		parts = path[2].split(".")
		if len(parts) != 4:
			raise RuntimeError("%s is not a valid table path" % str(path))
		block, division, _, _ = parts
		line = 1 + int(path[-1])  # index of line inside this block division
		grid = ".".join(map(str, (block, division, line, column)))
		return predictor, label, grid, str(0)

	def __call__(self, lines):
		if not self._columns:  # a page without tables: nothing to split
			return [(path, line, None) for path, line in lines.items()]
		line_parts = []
		for path, line in lines.items():
			line_columns = self._columns.get(path[:3])
			if line_columns is None:
				line_parts.append((path, line, None))
			else:
				line_columns = [None] + line_columns + [None]
				for i, (x0, x1) in enumerate(zip(line_columns, line_columns[1:])):
					line_parts.append((self._column_path(path, 1 + i), line, (x0, x1)))
		return line_parts


class _PathStrips(collections.abc.Sequence):
	"""[(line_path, strip)] over the packed strips of a page: the strips are already in host memory, the
	(ysteps, xsteps) view of one is only made when that item is read."""

	def __init__(self, paths, strips):
		self._paths, self._strips = paths, strips

	def __len__(self):
		return len(self._paths)

	def __getitem__(self, i):
		if isinstance(i, slice):
			return [self[k] for k in range(*i.indices(len(self)))]
		return self._paths[i], self._strips[i]

	def __iter__(self):
		return zip(self._paths, self._strips)


class LineExtractor:
	def __init__(self, tables, line_height, options, min_confidence=0.5):
		self._options = options
		self._line_height = line_height
		assert self._line_height is not None

		if self._options["binarize"].strip():
			self._binarizer = _binarize.from_string(self._options["binarize"])
		else:
			self._binarizer = None

		self._min_confidence = min_confidence
		self._rewriter = LineRewriter(tables)

	def _filtered_parts(self, lines, ignored=None):
		if ignored is not None:
			lines = dict(
				(k, v) for k, v in lines.items()
				if not ignored(tuple(k[:2])))

		min_confidence = self._min_confidence
		filtered_lines = dict()
		for path, line in lines.items():
			if line._confidence < min_confidence:
				logging.info("skipping line %s with confidence %.1f" % (
					str(path), line.confidence))
			else:
				filtered_lines[path] = line

		return self._rewriter(filtered_lines)

	def extract_arrays(self, lines, ignored=None):
		"""Sequence of (line_path, uint8 ndarray) -- the batched fast path; every strip is in host memory when
		this returns, the ndarray view of one is made when the item is read."""
		line_parts = self._filtered_parts(lines, ignored)
		if not line_parts:
			return []
		if self._options.get("do_not_dewarp"):
			# secondary path: per-line affine deskew (reference block.py:155-193)
			out = []
			for path, line, column in line_parts:
				im = line.image(target_height=self._line_height, column=column, dewarped=False,
					deskewed=not self._options.get("do_not_deskew"), binarizer=self._binarizer)
				out.append((path, im))
			return out
		dewarper = line_parts[0][1].block.page.dewarper
		if all(column is None for _, _, column in line_parts):
			# no table cells on this page: pack the descriptors in one vectorised step
			records = ops.pack_line_arrays(
				[line._p for _, line, _ in line_parts], [line._right for _, line, _ in line_parts],
				[line._up for _, line, _ in line_parts], ysteps=self._line_height)
		else:
			records = [line.record(ysteps=self._line_height, column=column) for _, line, column in line_parts]
		spec = self._binarizer.spec if self._binarizer is not None else None
		strips, _ = extract_strips(dewarper, records, binarize=spec)
		return _PathStrips([part[0] for part in line_parts], strips)

	def extract_device(self, lines, ignored=None, min_width=6, min_height=6, pad_value=0):
		"""The OCR stage's hand-off without the host round trip (reference origami/batch/detect/ocr.py:
		238-268): returns (kept_paths, batch, widths, empty_paths) where batch is a CUDA uint8 tensor
		(n_kept, line_height, w_max) holding the strips of at least min_width x min_height pixels in
		line order, right-padded with pad_value, widths a CUDA int32 tensor (n_kept) and empty_paths the
		lines the reference would write an empty transcription for.  The tensors export through
		DLPack (torch.utils.dlpack / tensor.__dlpack__) to whatever framework runs the recogniser."""
		import numpy as np
		import torch
		from ...core.block import _plan
		line_parts = self._filtered_parts(lines, ignored)
		if not line_parts:
			return [], None, None, []
		if self._options.get("do_not_dewarp"):
			raise ValueError("extract_device covers the dewarped path")
		dewarper = line_parts[0][1].block.page.dewarper
		grid = dewarper.grid
		records = [line.record(ysteps=self._line_height, column=column) for _, line, column in line_parts]
		plan, steps, offs = _plan(grid, records)
		dev = grid._device
		gray = torch.empty((max(int(offs[-1]), 1),), dtype=torch.uint8, device=dev)
		spec = self._binarizer.spec if self._binarizer is not None else None
		ops.extract_lines(dewarper.device_image, grid.device_grid, grid.resolution, plan, gray, binarize=spec,
			dense_map=grid._device_points)
		keep = (steps[:, 0] >= min_width) & (steps[:, 1] >= min_height)
		w_max = int(steps[keep, 0].max()) if keep.any() else 1
		w_max = (w_max + 3) // 4 * 4
		batch, widths, kept, n_kept = ops.pack_strips(gray, plan, self._line_height, w_max, min_width, min_height, pad_value)
		nk = int(keep.sum())
		kept_paths = [line_parts[i][0] for i in np.nonzero(keep)[0]]
		empty_paths = [line_parts[i][0] for i in np.nonzero(~keep)[0]]
		return kept_paths, batch[:nk], widths[:nk], empty_paths

	def __call__(self, lines, ignored=None):
		return [(path, im if isinstance(im, PIL.Image.Image) else PIL.Image.fromarray(im))
			for path, im in self.extract_arrays(lines, ignored)]
