"""
Seeded synthetic artifacts for the page-geometry path:
a broadsheet page, a sparse dewarping grid that overruns the page on all sides,
text lines in dewarped space, a label map and per-model class probabilities.

Pure NumPy; used by tests, bench.py and __graft_entry__.smoke().  The layouts follow
the reference's artifact formats: dewarp.zip (origami/core/dewarp.py:537-553),
lines.zip (docs/formats.md:223-246), segment.zip label maps (docs/formats.md:55-57).
"""

import math

import numpy as np


def grid_shape(width, height, cell=25):
	"""Sparse grid shape (gh, gw) that covers a width x height page with one extra cell
	on every side, as GridFactory's border extension does (origami/core/dewarp.py:341-419)."""
	return int(math.ceil(height / cell)) + 2, int(math.ceil(width / cell)) + 2


def make_page(width, height, seed=0, kind="text", dpi=600):
	"""8-bit gray page (height, width).  kind="text": paper 235+-6 with 6-8 columns of
	dark glyph-like blobs on a fixed line pitch; kind="noise": uniform noise (worst case
	for interpolation parity)."""
	rng = np.random.default_rng(1234 + seed)
	if kind == "noise":
		return rng.integers(0, 256, size=(height, width), dtype=np.uint8)
	page = rng.normal(235.0, 6.0, size=(height, width)).astype(np.float32)
	pitch = 28 if dpi >= 600 else 20
	glyph = 3 if dpi >= 600 else 2
	n_cols = int(rng.integers(6, 9))
	margin = width // 30
	col_w = (width - 2 * margin) // n_cols
	gh, gw = height // glyph, width // glyph
	ink = rng.random((gh, gw)) < 0.45
	rows = (np.arange(gh) * glyph) % pitch
	ink &= ((rows > pitch * 0.25) & (rows < pitch * 0.8))[:, None]
	cols = (np.arange(gw) * glyph - margin) % col_w
	ink &= ((cols > col_w * 0.04) & (cols < col_w * 0.96))[None, :]
	ink &= ((np.arange(gw) * glyph > margin) & (np.arange(gw) * glyph < width - margin))[None, :]
	ink = np.repeat(np.repeat(ink, glyph, axis=0), glyph, axis=1)
	dark = rng.normal(40.0, 25.0, size=ink.shape).astype(np.float32)
	sub = page[:ink.shape[0], :ink.shape[1]]
	sub[ink] = dark[ink]
	return np.clip(np.rint(page), 0, 255).astype(np.uint8)


def make_grid(width, height, cell=25, seed=0):
	"""Sparse grid (gh, gw, 2) fp32: node (i, j) of the dewarped lattice -> (x, y) on the
	warped page (docs/formats.md:298-306), a smooth sheet warp that leaves the page on all
	sides so border taps are exercised."""
	rng = np.random.default_rng(4321 + seed)
	gh, gw = grid_shape(width, height, cell)
	ax, ay = rng.uniform(4, 12, size=2)
	bx, by = rng.uniform(0, 20, size=2)
	ph1, ph2 = rng.uniform(0, 2 * np.pi, size=2)
	i = np.arange(gh, dtype=np.float64)[:, None]
	j = np.arange(gw, dtype=np.float64)[None, :]
	x = (j - 1) * cell + ax * np.sin(2 * np.pi * 1.5 * i / gh + ph1) + bx * (j / gw) ** 2
	y = (i - 1) * cell + ay * np.cos(2 * np.pi * 1.25 * j / gw + ph2) + by * (i / gh) ** 2
	return np.dstack([x, y]).astype(np.float32)


def make_lines(width_d, height_d, n_lines, seed=0, dpi=600, table_fraction=0.05):
	"""Text lines in DEWARPED space.  Returns a list of dicts with the lines.zip fields
	p (bottom-left), right, up (docs/formats.md:245-246; `up` points to smaller y),
	confidence, and an optional column (x0, x1) split for table rows
	(origami/batch/core/lines.py:68-82)."""
	if n_lines <= 0:
		return []
	rng = np.random.default_rng(9876 + seed)
	n_cols = 8
	col_w = width_d / n_cols
	scale = dpi / 300.0
	lines = []
	per_col = int(math.ceil(n_lines / n_cols))
	pitch = (height_d - 80 * scale) / per_col
	k = 0
	for c in range(n_cols):
		for r in range(per_col):
			if k >= n_lines:
				break
			theta = rng.uniform(-0.01, 0.01)
			h = rng.uniform(22, 34) * scale
			full = rng.random() < 0.7
			u = 1.0 if full else rng.uniform(0.35, 1.0)
			w = col_w * 0.92 * u
			x0 = c * col_w + col_w * 0.04 + rng.uniform(-2, 2)
			y0 = 40 * scale + (r + 1) * pitch + rng.uniform(-1, 1)
			line = dict(
				p=[float(x0), float(y0)],
				right=[float(w), float(w * math.tan(theta))],
				up=[float(-h * math.sin(theta)), float(-h * math.cos(theta))],
				confidence=float(rng.uniform(0.3, 1.0)) if rng.random() < 0.1 else 1.0,
				column=None)
			if rng.random() < table_fraction:
				n_split = int(rng.integers(2, 5))
				xs = np.sort(rng.uniform(x0 + 0.1 * w, x0 + 0.9 * w, size=n_split - 1))
				bounds = [None] + [float(v) for v in xs] + [None]
				line["splits"] = list(zip(bounds[:-1], bounds[1:]))
			lines.append(line)
			k += 1
	# a few exactly axis-aligned / integral lines: np.linspace takes its (i/div)*delta
	# branch when a step component is exactly zero, and integral sample coordinates hit
	# the cell-boundary case of the inverse lookup.
	for t in range(min(4, len(lines))):
		ln = lines[t]
		ln["p"] = [float(round(ln["p"][0])), float(round(ln["p"][1]))]
		ln["right"] = [float(round(ln["right"][0])), 0.0]
		ln["up"] = [0.0, float(-round(-ln["up"][1]))]
	return lines


def make_labels(width, height, n_classes=4, seed=0):
	"""Label map u8 (height, width) with smooth blobs (segment.zip payload)."""
	rng = np.random.default_rng(2468 + seed)
	yy = np.arange(height, dtype=np.float32)[:, None]
	xx = np.arange(width, dtype=np.float32)[None, :]
	score = np.zeros((n_classes, height, width), dtype=np.float32)
	for c in range(n_classes):
		fx, fy = rng.uniform(0.002, 0.02, size=2)
		px, py = rng.uniform(0, 6.28, size=2)
		score[c] = np.sin(xx * fx + px) * np.cos(yy * fy + py) + (0.4 if c == 0 else 0.0)
	return np.argmax(score, axis=0).astype(np.uint8)


def make_probabilities(n_models, width, height, n_classes=4, seed=0):
	"""Softmaxed class scores fp32 (M, height, width, C), one per ensemble member
	(origami/core/predict.py:281: model.predict(tile) outputs)."""
	rng = np.random.default_rng(1357 + seed)
	z = rng.standard_normal((n_models, height, width, n_classes)).astype(np.float32)
	z += rng.uniform(-0.5, 0.5, size=(1, 1, 1, n_classes)).astype(np.float32)
	z = np.exp(z - z.max(axis=-1, keepdims=True))
	return (z / z.sum(axis=-1, keepdims=True)).astype(np.float32)
