"""GPU parity of the contour transform: Grid.transformer and the batched
contour dewarping against SciPy's LinearNDInterpolator (the reference's own evaluator).  Floating
point: 1e-9 relative inside the hull of the lattice (same qhull triangulation on both sides), 1e-6
between the hull and the bounding box, where the extrapolation points come from shapely / sympy
geometry in the reference and are restated on both sides here (unpinned)."""

import numpy as np
import pytest

from oracle import exact
from origami_b200 import ops, synth
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


def _rings():
	rng = np.random.default_rng(7)
	star = []
	for k in range(14):
		r = 28 if k % 2 else 11
		star.append([45 + r * np.cos(k * np.pi / 7), 33 + r * np.sin(k * np.pi / 7)])
	return {
		"rectangle": np.array([[3, 4], [70, 4], [70, 30], [3, 30]], float),
		"thin sliver": np.array([[0, 0], [80, 1], [80, 3]], float),
		"1/510 sliver": np.array([[0, 10], [255, 11], [0, 11]], float),
		"convex": np.array([[10.4, 5.2], [60.7, 2.1], [85.2, 20.9], [70.3, 44.6], [20.5, 40.1], [2.2, 22.8]]),
		"star (concave)": np.array(star),
		"clockwise": np.array([[10, 10], [10, 40], [50, 40], [50, 10]], float),
		"leaves the raster": np.array([[-20, -10], [60, -15], [120, 30], [40, 80], [-30, 40]], float),
		"random": rng.uniform(0, 60, (9, 2)),
	}


def test_polygon_mask_matches_exact_coverage(cuda):
	from origami_b200 import ops
	w, h = 90, 64
	for name, ring in _rings().items():
		if name == "random":
			continue  # self-intersecting: winding rule territory, not produced by the line detector
		pts = np.round(ring).astype(np.int32)
		got = ops.polygon_mask(pts, w, h, device=cuda).cpu().numpy() > 0
		cov = exact.polygon_coverage(pts, w, h)
		want = cov >= 0.5 / 255.0
		near = np.abs(cov - 0.5 / 255.0) < 1e-12  # exactly on cairo's rounding step: either answer is defensible
		assert np.array_equal(got[~near], want[~near]), name
		assert got.sum() > 0


def test_mask_cutout_and_deskewed_line_with_polygon(cuda, setup):
	import cv2
	from origami_b200.core.mask import Mask
	W, H, s, g, grid, _ = setup
	page_px = synth.make_page(W, H, seed=4, kind="noise")
	ring = np.array([[100.3, 200.2], [420.9, 204.7], [430.0, 230.0], [260.5, 244.4], [99.1, 236.5]])
	mask = Mask(ring)
	x0, y0, x1, y1 = mask.bounds
	assert (x0, y0, x1, y1) == (99, 200, 430, 245)
	want = exact.polygon_mask(ring - np.array([x0, y0]), x1 - x0, y1 - y0)
	assert np.array_equal(mask.binary, want)
	cut = mask.cutout(page_px, background=7)
	ref = page_px[y0:y1, x0:x1].copy()
	ref[~want] = 7
	assert np.array_equal(cut, ref)
	# Line._extract_deskewed: warpAffine, then the cut-out through the affinely moved polygon
	wkt = "POLYGON ((" + ", ".join("%r %r" % (float(x), float(y)) for x, y in np.vstack([ring, ring[:1]])) + "))"
	block = Block(Page(page_px), wkt, Stage.WARPED)
	line = Line(block, [100.3, 236.2], [320.6, 4.5], [1.2, -30.0], wkt=wkt)
	got = line._extract_deskewed(page_px, 48, background=200)
	p, right, up = line._p, line._right, line._up
	width = int(np.ceil(np.linalg.norm(right)))
	M = cv2.getAffineTransform(np.array([p, p + right, p + up]).astype(np.float32), np.array([(0, 47), (width, 47), (0, 0)]).astype(np.float32))
	warped = exact.warp_affine_linear(page_px, M, width, 48, border=200)
	moved = ring @ M[:, :2].T + M[:, 2]
	inside = exact.polygon_mask(moved, width, 48)
	warped[~inside] = 200
	assert np.array_equal(got, warped)
	assert abs(block.background - np.quantile(page_px[200:245, 99:430], 0.8)) == 0  # the bounding box, as the reference
	assert line.warped_image.size == (331, 45)


@pytest.mark.parametrize("shape,dtype", [((301, 420), "u8"), ((256, 300), "f32"), ((1033, 200), "f32"), ((2, 40), "u8")])
def test_split_detector_matches_scipy(cuda, shape, dtype):
	"""layout.py:915-944 on a two-column text crop: column quantiles of the periodogram within float32
	round-off of SciPy's (complex64 pocketfft there, float64 DFT here), same peaks."""
	from origami_b200.batch.detect.layout import SplitDetector, crop
	h, w = shape
	page = synth.make_page(w + 40, h + 30, seed=h, kind="text", dpi=300)
	px, (x0, y0) = crop(page, (20.5, 10.2, 20.5 + w, 10.2 + h))
	assert px.shape == (h, w) and (x0, y0) == (20, 10)
	px = np.ascontiguousarray(px)
	px[:, w // 2 - 6:w // 2 + 6] = 235  # a whitespace gutter
	if dtype == "f32":
		px = (exact.binarize(px, ("otsu",)) > 0).astype(np.float32)  # what Regions.binarized hands over
	det = SplitDetector()
	peaks, info, colq = exact.split_detector(px, scale=20)
	got_q = det.column_quantiles(px, cuda)
	assert np.allclose(got_q, colq, rtol=2e-4, atol=1e-6 * max(float(colq.max()), 1e-12))
	if h > 16:
		gp, gi = det(px, 20, cuda)
		assert np.array_equal(gp, peaks) and np.allclose(gi["peak_heights"], info["peak_heights"], rtol=1e-3, atol=1e-6)


@pytest.mark.parametrize("transposed", [False, True])
def test_split_detector_on_all_regions_of_a_page(cuda, transposed):
	"""FixSpillOverH / FixSpillOverV (layout.py:947-1075) run the detector contour by contour on
	crop(regions.binarized, contour) (resp. its transpose); here all crops of the page go through one launch
	and must give what the per-crop path and the SciPy oracle give."""
	import torch
	from origami_b200.batch.detect.layout import SplitDetector, crop
	page = synth.make_page(900, 700, seed=3, kind="text", dpi=300)
	binar = (exact.binarize(page, ("otsu",)) > 0).astype(np.float32)
	bounds = [(20.5, 10.2, 420.5, 311.2), (-5.0, 300.0, 250.0, 690.0), (400.0, 350.0, 905.0, 705.0), (600.0, 20.0, 640.0, 60.0),
		(100.0, 100.0, 100.0, 300.0), (10.0, 500.0, 890.0, 540.0)]
	scales = [20, 18, 25, 6, 20, 12]
	det = SplitDetector()
	got = det.detect_regions(torch.from_numpy(binar).to(cuda), bounds, scales, transposed=transposed)
	assert len(got) == len(bounds)
	for (peaks, info, origin), b, sc in zip(got, bounds, scales):
		px, org = crop(binar, b)
		assert origin == org
		if px.size == 0:
			assert len(peaks) == 0
			continue
		px = np.ascontiguousarray(px.transpose() if transposed else px)
		want_p, want_i, colq = exact.split_detector(px, scale=sc)
		assert np.array_equal(peaks, want_p) and np.allclose(info["peak_heights"], want_i["peak_heights"], rtol=1e-3, atol=1e-6)
		one = det.column_quantiles(px, cuda)
		assert np.allclose(one, colq, rtol=2e-4, atol=1e-6 * max(float(colq.max()), 1e-12))
	with pytest.raises(ValueError):
		ops.region_psd_quantile(torch.from_numpy(binar).to(cuda), [(0, 0, 901, 10)])


def _flow_samples(W, H, n, seed):
	rng = np.random.default_rng(seed)
	ph = rng.uniform(0, [W, H], (n, 2))
	ah = 0.02 * np.sin(ph[:, 1] / 200) + 0.01 * np.cos(ph[:, 0] / 170)
	pv = rng.uniform(0, [W, H], (n, 2))
	av = np.pi / 2 + 0.02 * np.cos(pv[:, 0] / 150) - 0.01 * np.sin(pv[:, 1] / 230)
	return (ph, ah), (pv, av)


@pytest.mark.parametrize("W,H,n,threads", [(600, 800, 60, 2), (1000, 700, 150, 1), (500, 500, 2, 2)])
def test_grid_factory_matches_the_numpy_scipy_restatement(cuda, W, H, n, threads):
	"""flow samples -> sparse dewarping grid (origami/core/dewarp.py:292-493).
	Floating point: float32 positions, CUDA cosf / sinf against NumPy's, float64 segment crossings
	against GEOS-free NumPy ones -> 5e-3 px over the whole grid; the shape must be equal."""
	from origami_b200.core.gridfactory import GridFactory, Samples
	from origami_b200.core.math import Geometry
	(ph, ah), (pv, av) = _flow_samples(W, H, n, seed=W + n)
	sh, sv = Samples(Geometry(W, H)), Samples(Geometry(W, H))
	sh.extend(list(ph), list(ah))
	sv.extend(list(pv), list(av))
	fac = GridFactory((W, H), sh, sv, grid_res=25, num_threads=threads)
	got = fac.grid_hv
	want = exact.grid_factory((W, H), (ph, ah), (pv, av), 25, num_threads=threads)
	assert got.dtype == np.float32 and got.shape == want.shape
	assert np.abs(got - want).max() < 5e-3
	# the grid covers the page (border extension) and is usable as a dewarping grid
	assert got[:, 0, 0].max() <= 2 and got[:, -1, 0].min() >= W - 2 and got[0, :, 1].max() <= 2 and got[-1, :, 1].min() >= H - 2  # (rows added last may lean back in a little)
	grid = Grid.create((W, H), sh, sv, grid_res=25, num_threads=threads)
	assert grid.resolution == 25 and np.array_equal(grid.points("sample"), got)
	page = synth.make_page(W, H, seed=1, kind="noise")
	dew = np.array(Page(page, grid).dewarped)
	assert np.array_equal(dew, exact.remap_linear(page, exact.expand_map(got, 25)))
