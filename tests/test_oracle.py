"""
CPU tests that pin the oracle (oracle/exact.py):
  * against the golden vectors in tests/golden/, produced by the unmodified reference;
  * against the libraries the reference calls (OpenCV generic path, SciPy, NumPy);
  * against independent second formulations for the two scikit-image thresholds
    (scikit-image itself is not installed: those two are "parity unpinned" vs skimage).
"""

import glob
import itertools
import os

import cv2
import numpy as np
import pytest
import scipy.interpolate

from oracle import exact, ref_import, reference_path
from origami_b200 import synth

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ref_*.npz")))


@pytest.fixture(autouse=True)
def _generic_opencv():
	cv2.setUseOptimized(False)
	yield
	cv2.setUseOptimized(True)


def _bits(a):
	return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
	z = np.load(path)
	page, grid, cell = z["page"], z["grid"], int(z["cell"])
	dense = exact.expand_map(grid, cell)
	assert np.array_equal(_bits(dense), _bits(z["dense"]))
	assert np.array_equal(exact.remap_linear(page, dense), z["dewarped"])
	size = (page.shape[1], page.shape[0])
	lab60 = z["labels"] * 60
	assert np.array_equal(exact.dewarp_image(lab60, size, grid, cell, exact.INTER_NEAREST), z["labels_nearest"])
	assert np.array_equal(exact.dewarp_image(lab60, size, grid, cell, exact.INTER_LINEAR), z["labels_linear"])
	assert np.array_equal(exact.inverse_yx(dense, z["inv_pts"]), z["inv_out"])
	for i in range(int(z["n_lines"])):
		p, right, up = z["line%d_pru" % i]
		for h in (48, 11):
			want = z["line%d_strip%d" % (i, h)]
			got = exact.dewarped_line_image(page, dense, p, right, up, h)
			assert got.shape == want.shape and np.array_equal(got, want), (i, h)
		half = exact.warped_grid(dense, p, right, up, xres=0.5, yres=0.5)
		assert half.shape == z["line%d_half" % i].shape
		assert np.array_equal(_bits(half), _bits(z["line%d_half" % i]))
		path_xy, height = exact.warped_path(dense, p, right, up, 1)
		assert np.array_equal(path_xy, z["line%d_path" % i]) and height == float(z["line%d_pathh" % i])
		counts = exact.sample_label_counts(z["labels"], page.shape, dense, p, right, up, 4)
		assert np.array_equal(counts, z["line%d_counts" % i])


def test_golden_files_present():
	assert len(GOLDEN) >= 3


@pytest.mark.skipif(not ref_import.available(), reason="/root/reference only exists in the build container")
def test_oracle_matches_live_reference():
	import PIL.Image
	Grid, Dewarper, Line = ref_import.load()
	W, H, s = 420, 300, 25
	page = synth.make_page(W, H, seed=31, kind="noise")
	g = synth.make_grid(W, H, s, seed=31)
	grid = Grid(g.copy(), s)
	pil = PIL.Image.fromarray(page)
	dw = Dewarper(pil, grid)
	dense = exact.expand_map(g, s)
	assert np.array_equal(_bits(grid.points("full")), _bits(dense))
	assert np.array_equal(np.array(dw.dewarped), exact.remap_linear(page, dense))
	block = ref_import._Block(ref_import._Page(pil, dw))
	for ln in synth.make_lines(dense.shape[1], dense.shape[0], 24, seed=31, dpi=150):
		L = ref_import.make_line(Line, block, ln["p"], ln["right"], ln["up"])
		assert np.array_equal(np.array(L.dewarped_image(48)),
			exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48))


@pytest.mark.parametrize("gh,gw,s", [(30, 20, 25), (31, 17, 8), (20, 33, 3), (2, 2, 5), (9, 7, 1), (12, 10, 2)])
def test_expand_plane_is_cv2_resize_generic(gh, gw, s):
	rng = np.random.default_rng(gh * 100 + gw)
	S = (rng.standard_normal((gh, gw)) * 50 + np.arange(gw)[None, :] * s + 4000).astype(np.float32)
	want = cv2.resize(S, (gw * s, gh * s), interpolation=cv2.INTER_LINEAR)
	assert np.array_equal(_bits(exact.expand_plane(S, s)), _bits(want))


def test_expand_map_close_to_default_opencv_path():
	"""The IPP-dispatched default path is not bit-reproducible; it stays within a few ulp."""
	g = synth.make_grid(2400, 3600, 25, seed=1)
	cv2.setUseOptimized(True)
	ref = reference_path.RefGrid(g, 25).points_full()
	mine = exact.expand_map(g, 25)
	assert np.abs(ref - mine).max() <= 4 * np.spacing(np.float32(4096))


def test_remap_linear_is_cv2_remap():
	rng = np.random.default_rng(5)
	src = rng.integers(0, 256, (211, 173), dtype=np.uint8)
	m = np.stack([rng.uniform(-6, 180, (300, 257)), rng.uniform(-6, 218, (300, 257))], axis=-1).astype(np.float32)
	m[:50, :50] = np.rint(m[:50, :50] * 32) / 32 + 1 / 64  # exact half-way cases of the 1/32 quantiser
	assert np.array_equal(exact.remap_linear(src, m), cv2.remap(src, m, None, cv2.INTER_LINEAR))


def test_remap_nearest_is_cv2_remap():
	rng = np.random.default_rng(6)
	src = rng.integers(0, 256, (90, 120), dtype=np.uint8)
	m = np.stack([rng.uniform(-4, 124, (200, 150)), rng.uniform(-4, 94, (200, 150))], axis=-1).astype(np.float32)
	m[:40, :40] = np.rint(m[:40, :40]) + 0.5  # ties: round half to even
	assert np.array_equal(exact.remap_nearest(src, m), cv2.remap(src, m, None, cv2.INTER_NEAREST))


def test_inverse_yx_is_scipy_rgi():
	rng = np.random.default_rng(7)
	dense = (rng.standard_normal((60, 45, 2)) * 20 + 3000).astype(np.float32)
	rgi = scipy.interpolate.RegularGridInterpolator((np.arange(60), np.arange(45)), dense,
		method="linear", bounds_error=False, fill_value=None)
	pts = np.stack([rng.uniform(-20, 80, 5000), rng.uniform(-20, 65, 5000)], axis=1)
	pts[:300] = np.rint(pts[:300])
	assert np.array_equal(exact.inverse_yx(dense, pts), rgi(pts))


@pytest.mark.parametrize("stop,num", [((3.7, -41.2), 48), ((250.0, 0.0), 33), ((0.0, -30.0), 2), ((0.0, 0.0), 5),
	((1e-3, 7.0), 1), ((-12.5, 0.3), 3)])
def test_linspace_restatement(stop, num):
	assert np.array_equal(exact._linspace2(stop, num), np.linspace([0, 0], stop, num))


def test_warped_grid_and_steps_match_reference_path():
	W, H, s = 500, 400, 25
	g = synth.make_grid(W, H, s, seed=3)
	dense = exact.expand_map(g, s)
	rg = reference_path.RefGrid(g, s)
	page = synth.make_page(W, H, seed=3, kind="noise")
	import PIL.Image
	pil = PIL.Image.fromarray(page)
	for ln in synth.make_lines(dense.shape[1], dense.shape[0], 30, seed=3, dpi=150):
		for kw in (dict(ysteps=48), dict(xres=0.5, yres=0.5), dict(ysteps=2, xres=2)):
			a = exact.warped_grid(dense, ln["p"], ln["right"], ln["up"], **kw)
			b = reference_path.warped_grid(rg, ln["p"], ln["right"], ln["up"], **kw)
			assert a.shape == b.shape and np.array_equal(_bits(a), _bits(b))
		col = (ln["p"][0] + 0.25 * ln["right"][0], ln["p"][0] + 0.75 * ln["right"][0])
		a = exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48, column=col)
		b = reference_path.line_image(pil, rg, ln["p"], ln["right"], ln["up"], 48, column=col)
		assert a.shape == b.shape and np.array_equal(a, b)


def test_reflect_index_is_np_pad():
	for n, pad in itertools.product((1, 2, 3, 5, 9), (1, 4, 11)):
		a = np.arange(n)
		want = np.pad(a, (pad, pad), mode="reflect")
		assert np.array_equal(exact._reflect_index(np.arange(-pad, n + pad), n), want)


def _sauvola_skimage_formulation(image, w, k=0.2, r=127.5):
	"""skimage 0.16.2 _mean_std written out with np.pad + float integral images."""
	left, right = w // 2 + 1, w // 2
	padded = np.pad(image.astype("float"), (left, right), mode="reflect")
	integral = np.cumsum(np.cumsum(padded, 0), 1)
	integral_sq = np.cumsum(np.cumsum(padded * padded, 0), 1)

	def corner(I):
		# ndi.correlate with the (w+1)^2 corner kernel, origin (w+1)//2, then crop
		h, wd = image.shape
		c = (w + 1) // 2
		out = np.empty(image.shape)
		for y in range(h):
			for x in range(wd):
				i, j = y + left, x + left
				out[y, x] = I[i - c, j - c] - I[i - c, j + w - c] - I[i + w - c, j - c] + I[i + w - c, j + w - c]
		return out

	m = corner(integral) / (w ** 2)
	g2 = corner(integral_sq) / (w ** 2)
	s = np.sqrt(np.clip(g2 - m * m, 0, None))
	return m * (1 + k * ((s / r) - 1))


@pytest.mark.parametrize("shape,w", [((23, 31), 15), ((12, 40), 3), ((9, 9), 7), ((5, 20), 9)])
def test_sauvola_two_formulations_agree(shape, w):
	rng = np.random.default_rng(w)
	img = rng.integers(0, 256, shape, dtype=np.uint8)
	assert np.array_equal(exact.threshold_sauvola(img, w), _sauvola_skimage_formulation(img, w))


def test_sauvola_window_validation():
	with pytest.raises(ValueError):
		exact.threshold_sauvola(np.zeros((8, 8), np.uint8), 4)
	with pytest.raises(ValueError):
		exact.threshold_sauvola(np.zeros((8, 8), np.uint8), 1)


def _otsu_skimage_formulation(image):
	"""skimage 0.16.2 threshold_otsu via np.histogram-style bins on integer data."""
	if image.min() == image.max():
		raise ValueError
	lo, hi = int(image.min()), int(image.max())
	hist = np.array([(image == v).sum() for v in range(lo, hi + 1)]).astype(float)
	bin_centers = np.arange(lo, hi + 1)
	weight1 = np.cumsum(hist)
	weight2 = np.cumsum(hist[::-1])[::-1]
	mean1 = np.cumsum(hist * bin_centers) / weight1
	mean2 = (np.cumsum((hist * bin_centers)[::-1]) / weight2[::-1])[::-1]
	variance12 = weight1[:-1] * weight2[1:] * (mean1[:-1] - mean2[1:]) ** 2
	return bin_centers[:-1][np.argmax(variance12)]


def test_otsu_two_formulations_agree():
	for seed in range(6):
		img = synth.make_page(160, 90, seed=seed, kind="text" if seed % 2 else "noise")
		assert exact.threshold_otsu(img) == _otsu_skimage_formulation(img)
	with pytest.raises(ValueError):
		exact.threshold_otsu(np.full((4, 4), 9, np.uint8))
	assert (exact.binarize(np.full((4, 4), 200, np.uint8), ("otsu",)) == 255).all()  # t = 128 fallback


def test_argmax_is_numpy():
	for M in (1, 2, 5):
		pr = synth.make_probabilities(M, 37, 29, 4, seed=M)
		want = np.argmax(np.sum([p[None] for p in pr], axis=0).squeeze(), axis=-1).astype(np.uint8)
		assert np.array_equal(exact.argmax_labels(pr), want)
		assert np.array_equal(reference_path.argmax_labels([p[None] for p in pr]), want)


def test_warp_affine_is_cv2():
	rng = np.random.default_rng(9)
	page = rng.integers(0, 256, (300, 400), dtype=np.uint8)
	for k in range(5):
		p = rng.uniform(40, 200, 2)
		right = np.array([rng.uniform(80, 180), rng.uniform(-6, 6)])
		up = np.array([rng.uniform(-3, 3), -rng.uniform(15, 30)])
		width = int(np.ceil(np.linalg.norm(right)))
		src_tri = np.array([p, p + right, p + up]).astype(np.float32)
		dst_tri = np.array([(0, 47), (width, 47), (0, 0)]).astype(np.float32)
		M = cv2.getAffineTransform(src_tri, dst_tri)
		assert np.allclose(exact.affine_from_points(src_tri, dst_tri), M, atol=1e-9)
		for flag in (cv2.INTER_LINEAR, cv2.INTER_AREA):
			want = cv2.warpAffine(page, M, (width, 48), flags=flag, borderMode=cv2.BORDER_CONSTANT, borderValue=200)
			assert np.array_equal(exact.warp_affine_linear(page, M, width, 48, border=200), want)


def test_grid_bbox_convention():
	g = np.array([[[1.2, 5.0], [7.9, -0.5]]], dtype=np.float32)
	assert exact.grid_bbox(g) == (1, -1, 8, 5)


def test_dilate3_is_3x3_maximum_with_zero_border():
	rng = np.random.default_rng(2)
	m = rng.random((23, 31)) < 0.05
	want = cv2.dilate(m.astype(np.uint8), np.ones((3, 3), np.uint8), borderType=cv2.BORDER_CONSTANT, borderValue=0) > 0
	assert np.array_equal(exact.dilate3(m), want)


def test_otsu_agrees_with_opencv():
	"""cv2.threshold(THRESH_OTSU) maximises the same between-class variance over the same 256 bins
	(another code base, float arithmetic of its own): a transcription slip in exact.threshold_otsu
	would show up here.  The two may differ only where the criterion is tied to within rounding."""
	for seed in range(12):
		img = synth.make_page(320, 200, seed=seed, kind="text" if seed % 2 else "noise", dpi=300)
		if seed % 3 == 0:
			img = (img // 3 + 40).astype(np.uint8)  # a narrow histogram
		t_cv, _ = cv2.threshold(img, 0, 255, cv2.THRESH_BINARY + cv2.THRESH_OTSU)
		t = exact.threshold_otsu(img)
		if int(t_cv) != t:
			hist = np.bincount(img.ravel(), minlength=256).astype(np.float64)

			def crit(k):
				w1, w2 = hist[:k + 1].sum(), hist[k + 1:].sum()
				m1 = (hist[:k + 1] * np.arange(k + 1)).sum() / w1
				m2 = (hist[k + 1:] * np.arange(k + 1, 256)).sum() / w2
				return w1 * w2 * (m1 - m2) ** 2
			assert abs(crit(int(t_cv)) - crit(t)) <= 1e-9 * crit(t), (seed, t_cv, t)
