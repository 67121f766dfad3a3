"""
LineRewriter and LineExtractor with the interface of the reference's
origami/batch/core/lines.py:51-149.  The reference maps Line.image over a ThreadPool(8);
here every line of the page goes through one batched kernel launch.
"""

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
		"""[(line_path, uint8 ndarray)] -- the batched fast path."""
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
		return [(part[0], strip) for part, strip in zip(line_parts, strips)]

	def __call__(self, lines, ignored=None):
		return [(path, im if isinstance(im, PIL.Image.Image) else PIL.Image.fromarray(im))
			for path, im in self.extract_arrays(lines, ignored)]
