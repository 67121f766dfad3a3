"""GPU parity of the contour transform: Grid.transformer and the batched
contour dewarping against SciPy's LinearNDInterpolator (the reference's own evaluator).  Floating
point: 1e-9 relative inside the hull of the lattice (same qhull triangulation on both sides), 1e-6
between the hull and the bounding box, where the extrapolation points come from shapely / sympy
geometry in the reference and are restated on both sides here (unpinned)."""

import numpy as np
import pytest

from oracle import exact
from origami_b200 import synth
from origami_b200.batch.detect import dewarp as bdewarp
from origami_b200.core.block import Block, Line, Stage
from origami_b200.core.dewarp import Grid
from origami_b200.core.page import Page

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup(cuda):
	W, H, s = 1500, 2000, 25
	g = synth.make_grid(W, H, s, seed=4)
	grid = Grid(g, s)
	return W, H, s, g, grid, exact.transformer(exact.expand_map(g, s), s)


def test_transformer_matches_scipy_on_contour_vertices(setup):
	W, H, s, g, grid, want_f = setup
	rng = np.random.default_rng(1)
	pts = np.stack([rng.uniform(-30, W + 30, 100000), rng.uniform(-30, H + 30, 100000)], axis=1)  # 10^5 vertices, some off the lattice
	got = grid.transformer.transform_points(pts)
	want, outside = want_f(pts)
	assert outside.sum() < 0.2 * len(pts) and (~outside).sum() > 50000
	scale = np.maximum(np.abs(want), 1.0)
	assert (np.abs(got - want)[~outside] <= 1e-9 * scale[~outside]).all()
	assert (np.abs(got - want)[outside] <= 1e-6 * scale[outside]).all()
	x, y = grid.transformer(pts[:50, 0], pts[:50, 1])
	assert np.array_equal(np.stack([x, y], axis=1), got[:50])
	# the transform inverts the grid: the warped position of sparse node (i, j) maps back to the centre of
	# its cell on the dewarped page (cv2.resize aligns pixel centres: dense pixel (j + 0.5) * s - 0.5)
	back = grid.transformer.transform_points(g[3:-3, 3:-3].reshape(-1, 2).astype(np.float64))
	jj, ii = np.meshgrid(np.arange(3, g.shape[1] - 3), np.arange(3, g.shape[0] - 3))
	assert np.abs(back - np.stack([(jj.ravel() + 0.5) * s - 0.5, (ii.ravel() + 0.5) * s - 0.5], axis=1)).max() < 0.1


def test_dewarped_contours_and_line_height(setup):
	W, H, s, g, grid, want_f = setup
	rng = np.random.default_rng(2)
	named = []
	for k in range(40):
		cx, cy = rng.uniform(200, W - 200), rng.uniform(200, H - 200)
		ang = np.sort(rng.uniform(0, 2 * np.pi, 12))
		r = rng.uniform(5 if k % 10 == 0 else 60, 8 if k % 10 == 0 else 150)
		ring = np.stack([cx + r * np.cos(ang), cy + r * np.sin(ang)], axis=1)
		ring = np.vstack([ring, ring[:1]])
		named.append(("regions/TEXT/%d.wkt" % k, bdewarp.format_wkt("POLYGON", [ring]).encode("utf8")))
	named.append(("separators/H/0.wkt", b"LINESTRING (100 300, 900 310, 1400 305)"))
	named.append(("meta.json", b"{}"))
	out = dict(bdewarp.dewarped_contours(named, grid.transformer, {"regions": 500.0}))
	assert "separators/H/0.wkt" in out and "meta.json" not in out
	lost = [n for n, _ in named if n.endswith(".wkt") and n not in out]
	assert lost == ["regions/TEXT/%d.wkt" % k for k in (0, 10, 20, 30)]  # the small ones fall below the minimum area
	for name, data in named:
		if name in out:
			kind, rings = bdewarp.parse_wkt(data.decode("utf8"))
			kind2, moved = bdewarp.parse_wkt(out[name].decode("utf8"))
			want, _ = want_f(rings[0])
			assert kind == kind2 and np.abs(moved[0] - want).max() <= 1e-8
	page = Page(synth.make_page(W, H, seed=4, kind="noise"), grid)
	line = Line(Block(page, None, Stage.WARPED), [400.0, 500.0], [300.0, 2.0], [1.0, -30.0])
	q, _ = want_f(np.array([[400.0, 500.0], [401.0, 470.0]]))
	assert abs(line.dewarped_height(page.dewarper) - np.linalg.norm(q[1] - q[0])) < 1e-8
