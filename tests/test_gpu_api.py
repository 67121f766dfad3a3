"""GPU parity of the host-side mirror of the reference interface (Grid, Dewarper, Page, Line,
LineExtractor, ConfidenceSampler, binarizers, argmax) against the reference path restated on
OpenCV / SciPy (oracle/reference_path.py, generic OpenCV code path) and the exact oracle."""

import enum
import types

import cv2
import numpy as np
import PIL.Image
import pytest
import torch

from oracle import exact, reference_path
from origami_b200 import ops, synth
from origami_b200.batch.core.lines import LineExtractor
from origami_b200.batch.detect.lines import ConfidenceSampler, scale_grid
from origami_b200.core import binarize, predict
from origami_b200.core.block import Block, Line, Stage
from origami_b200.core.dewarp import Dewarper, Grid, INTER_NEAREST
from origami_b200.core.page import Page

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _generic_opencv():
	cv2.setUseOptimized(False)
	yield
	cv2.setUseOptimized(True)


@pytest.fixture(scope="module")
def small(cuda):
	W, H, s = 700, 500, 25
	page = synth.make_page(W, H, seed=12, kind="text", dpi=300)
	g = synth.make_grid(W, H, s, seed=12)
	grid = Grid(g, s)
	pg = Page(page, grid)
	ref = reference_path.RefGrid(g, s)
	lines = synth.make_lines(g.shape[1] * s, g.shape[0] * s, 64, seed=12, dpi=300, table_fraction=0.3)
	return types.SimpleNamespace(page=page, g=g, s=s, grid=grid, pg=pg, ref=ref, lines=lines,
		pil=PIL.Image.fromarray(page))


def test_grid_points_and_dewarper(small):
	full = small.grid.points("full")
	want = small.ref.points_full()
	assert full.dtype == np.float32 and np.array_equal(full.view(np.uint32), want.view(np.uint32))
	assert small.grid.points("full") is full  # cached like the reference's lru_cache
	dw = small.pg.dewarped
	assert isinstance(dw, PIL.Image.Image) and dw.size == small.grid.geometry.size
	assert np.array_equal(np.array(dw), np.array(reference_path.dewarped(small.pil, small.ref)))
	assert small.pg.size(True) == dw.size and small.pg.pixels(False).shape == small.page.shape


def test_dewarp_image_does_not_corrupt_cached_map(small):
	labels = synth.make_labels(350, 250, seed=4) * 60
	lab_pil = PIL.Image.fromarray(labels)
	before = small.grid.points("full").copy()
	got = small.pg.dewarper.dewarp_image(lab_pil, INTER_NEAREST)
	want = reference_path.dewarp_image(small.pil, small.ref, lab_pil, cv2.INTER_NEAREST)
	assert np.array_equal(np.array(got), np.array(want))
	assert np.array_equal(small.grid.points("full"), before)  # the reference halves its cached map here


def test_inverse_is_scipy(small):
	rng = np.random.default_rng(3)
	Hd, Wd = small.grid.points("full").shape[:2]
	pts_xy = np.stack([rng.uniform(-20, Wd + 20, 3000), rng.uniform(-20, Hd + 20, 3000)], axis=1)
	got = small.grid.inverse(pts_xy)
	want = small.ref.inverse_yx()(np.flip(pts_xy, axis=-1))
	assert np.array_equal(got, want)


def test_line_methods(small):
	block = Block(small.pg, None, Stage.RELIABLE)
	for ln in small.lines[:16]:
		line = Line(block, ln["p"], ln["right"], ln["up"])
		wg = line.warped_grid(xres=0.5, yres=0.5)
		ref = reference_path.warped_grid(small.ref, ln["p"], ln["right"], ln["up"], xres=0.5, yres=0.5)
		assert wg.shape == ref.shape and np.array_equal(wg.view(np.uint32), ref.view(np.uint32))
		im = line.dewarped_image(48)
		want = reference_path.line_image(small.pil, small.ref, ln["p"], ln["right"], ln["up"], 48)
		assert np.array_equal(np.array(im), want)
		path, height = line.warped_path(1)
		p2, h2 = exact.warped_path(small.ref.points_full(), ln["p"], ln["right"], ln["up"], 1)
		assert np.array_equal(path, p2) and height == h2
		bim = line.image(target_height=32, binarizer=binarize.otsu())
		want = reference_path.line_image(small.pil, small.ref, ln["p"], ln["right"], ln["up"], 32, binarize=("otsu",))
		assert np.array_equal(np.array(bim), want)


@pytest.mark.parametrize("spec", ["", "otsu", "sauvola(window_size=15)"])
def test_line_extractor_matches_reference_loop(small, spec):
	block = Block(small.pg, None, Stage.RELIABLE)
	lines, want_parts = dict(), []
	columns = dict()
	for i, ln in enumerate(small.lines):
		if "splits" in ln:
			path = ("regions", "TABULAR", "%d.1.1.1" % i, "0")
			xs = [b for (a, b) in ln["splits"][:-1]]
			columns["/".join(path[:3])] = xs
			for col in ln["splits"]:
				want_parts.append(dict(p=ln["p"], right=ln["right"], up=ln["up"], column=col))
		else:
			path = ("regions", "TEXT", str(i), "0")
			if ln["confidence"] >= 0.5:
				want_parts.append(dict(p=ln["p"], right=ln["right"], up=ln["up"], column=None))
		lines[path] = Line(block, ln["p"], ln["right"], ln["up"], confidence=ln["confidence"] if "splits" not in ln else 1.0)
	ex = LineExtractor(dict(columns=columns), 48, dict(binarize=spec, do_not_dewarp=False, do_not_deskew=False), min_confidence=0.5)
	got = ex(lines)
	bspec = None if not spec else (("otsu",) if spec == "otsu" else ("sauvola", 15))
	want = reference_path.extract_lines(small.pil, small.ref, want_parts, 48, bspec)
	assert len(got) == len(want) > 40
	for (path, im), w in zip(got, want):
		assert isinstance(im, PIL.Image.Image)
		assert np.array_equal(np.array(im), w), path


def test_confidence_sampler(small):
	class Label(enum.Enum):
		BACKGROUND = 0
		TEXT = 1
		TABULAR = 2
		ILLUSTRATION = 3

	labels = synth.make_labels(320, 240, seed=6)
	pred = types.SimpleNamespace(name="regions", labels=labels, classes=Label)
	seg = types.SimpleNamespace(predictions=[pred])
	block = Block(small.pg, None, Stage.RELIABLE)
	sampler = ConfidenceSampler({("regions", "TEXT", "0"): block}, seg)
	items = []
	for i, ln in enumerate(small.lines[:24]):
		items.append((("regions", "TEXT", str(i), "0"), Line(block, ln["p"], ln["right"], ln["up"])))
	batched = sampler.sample_all(items)
	for (path, line), ev, ln in zip(items, batched, small.lines):
		counts = reference_path.confidence_counts(small.ref, labels, small.page.shape, ln["p"], ln["right"], ln["up"], 4)
		want = dict(("regions/%s" % k.name, counts[k.value] / counts.sum()) for k in Label)
		assert ev == want
	assert sampler(*items[0]) == batched[0]
	g = np.ones((2, 2, 2), np.float32)
	scale_grid((100, 200), (50, 50), g)
	assert g[0, 0].tolist() == [0.25, 0.5]


def test_page_binarized_and_binarizers(small):
	want = exact.binarize(small.page, ("sauvola", 15))
	assert np.array_equal(np.array(small.pg.binarized), want)
	assert np.array_equal(np.array(binarize.from_string("otsu")(small.pil)), exact.binarize(small.page, ("otsu",)))


def test_predict_argmax_and_masks(cuda):
	pr = synth.make_probabilities(5, 64, 48, 4, seed=3)
	labels = predict.argmax_labels([p[None] for p in pr])
	assert np.array_equal(labels, reference_path.argmax_labels([p[None] for p in pr]))

	class C(enum.Enum):
		TEXT = 1
		TABULAR = 2

	masks = predict.class_masks(labels, C)
	for c in C:
		assert masks[c].dtype == bool and np.array_equal(masks[c], labels == c.value)


def test_deskewed_image_warp_stage(small):
	block = Block(small.pg, None, Stage.WARPED)
	ln = small.lines[5]
	line = Line(block, ln["p"], ln["right"], ln["up"])
	p, right, up = (np.array(ln[k]) for k in ("p", "right", "up"))
	width = int(np.ceil(np.linalg.norm(right)))
	M = cv2.getAffineTransform(np.array([p, p + right, p + up]).astype(np.float32),
		np.array([(0, 47), (width, 47), (0, 0)]).astype(np.float32))
	got = line._extract_deskewed(small.page, 48, background=231)
	Mm = exact.affine_from_points(np.array([p, p + right, p + up]), np.array([(0, 47), (width, 47), (0, 0)]))
	want = cv2.warpAffine(small.page, Mm, (width, 48), flags=cv2.INTER_AREA, borderMode=cv2.BORDER_CONSTANT, borderValue=231)
	assert np.array_equal(got, want)
	assert np.allclose(M, Mm, atol=1e-9)


def test_full_size_page_against_opencv(cuda):
	"""BASELINE configs[1] size: 4800x7200 page, map + dewarped page bit-equal to OpenCV's generic
	path, strips of 200 lines bit-equal to the oracle, and size-independent properties."""
	W, H, s = 4800, 7200, 25
	page = synth.make_page(W, H, seed=3, kind="noise")
	g = synth.make_grid(W, H, s, seed=3)
	ref = reference_path.RefGrid(g, s)
	want_map = ref.points_full()
	gt = torch.from_numpy(g).to(cuda)
	src = ops.to_device_image(page, cuda)
	dmap = torch.empty(want_map.shape, dtype=torch.float32, device=cuda)
	dst, _ = ops.remap_page(src, gt, s, map_out=dmap)
	assert np.array_equal(dmap.cpu().numpy().view(np.uint32), want_map.view(np.uint32))
	want_img = cv2.remap(page, want_map, None, interpolation=cv2.INTER_LINEAR)
	assert np.array_equal(dst.cpu().numpy(), want_img)
	# property: remapping a constant page gives the constant wherever all four taps are inside
	const = ops.to_device_image(np.full((H, W), 77, np.uint8), cuda)
	out, _ = ops.remap_page(const, gt, s)
	out = out.cpu().numpy()
	inside = (want_map[..., 0] >= 0) & (want_map[..., 0] < W - 1) & (want_map[..., 1] >= 0) & (want_map[..., 1] < H - 1)
	assert (out[inside] == 77).all() and out.max() <= 77
	lines = synth.make_lines(want_map.shape[1], want_map.shape[0], 2000, seed=3, dpi=600)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines])
	plan = ops.line_steps(gt, s, ops.lines_to_device(la, cuda), len(lines))
	offs = plan.pix_offsets.cpu().numpy()
	a = torch.empty((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	b = torch.empty_like(a)
	ops.extract_lines(src, gt, s, plan, a)
	ops.extract_lines(src, gt, s, plan, b, dense_map=dmap, map_gather=True)
	assert torch.equal(a, b)  # line-owning and map-gather engines agree on all 2000 lines
	a = a.cpu().numpy()
	for i in range(0, 2000, 10):
		l = lines[i]
		want = exact.dewarped_line_image(page, want_map, l["p"], l["right"], l["up"], 48)
		assert np.array_equal(a[offs[i]:offs[i + 1]].reshape(want.shape), want), i
	# property: Otsu output is idempotent
	o1 = ops.binarize_strips(torch.from_numpy(a).to(cuda), plan, ops.OG_BIN_OTSU)
	o2 = ops.binarize_strips(o1, plan, ops.OG_BIN_OTSU)
	assert torch.equal(o1, o2)


def test_regions_binarized(small):
	from origami_b200.batch.detect.layout import regions_binarized, sauvola_window
	assert [sauvola_window(v) for v in (4, 20, 31, 60)] == [3, 11, 15, 31]
	labels = (synth.make_labels(350, 250, seed=9) == 2).astype(np.uint8) * 3  # sparse "separator" pixels, bg = 0
	got = regions_binarized(small.pg, 31, [(labels, 0)])
	want = exact.regions_binarized(small.page, small.g, small.s, 31, [(labels, 0)])
	assert got.dtype == np.float32 and np.array_equal(got, want)
	assert np.array_equal(regions_binarized(small.pg, 20), exact.regions_binarized(small.page, small.g, small.s, 20))
	dev = regions_binarized(small.pg, 31, [(labels, 0)], on_device=True)  # stays on the GPU for SplitDetector.detect_regions
	assert dev.is_cuda and dev.dtype == torch.float32 and np.array_equal(dev.cpu().numpy(), want)


@pytest.mark.parametrize("spec", ["", "otsu"])
def test_page_pipeline_matches_per_page_api(cuda, spec):
	"""The multi-page pipeline (pinned staging, two streams) returns the same bytes as the
	per-page objects, in job order, for pages of different sizes."""
	from origami_b200.batch.core.pipeline import PagePipeline, pinned_page
	jobs, want = [], []
	for k, (W, H, n) in enumerate([(700, 500, 40), (333, 421, 25), (900, 300, 30), (512, 512, 0), (640, 480, 33)]):
		page = synth.make_page(W, H, seed=50 + k, kind="noise")
		g = synth.make_grid(W, H, 25, seed=50 + k)
		lines = synth.make_lines(g.shape[1] * 25, g.shape[0] * 25, n, seed=50 + k, dpi=300, table_fraction=0.0)
		arr = np.array([[l["p"], l["right"], l["up"]] for l in lines], dtype=np.float64).reshape(-1, 3, 2)
		jobs.append((pinned_page(page) if k % 2 else page, g, 25, arr))  # pinned and pageable inputs
		dense = exact.expand_map(g, 25)
		strips = [exact.binarize(exact.dewarped_line_image(page, dense, l["p"], l["right"], l["up"], 48),
			("otsu",) if spec else None) for l in lines]
		want.append((exact.remap_linear(page, dense), strips))
	pipe = PagePipeline(line_height=48, binarize=spec)
	n_seen = 0
	for res, (wd, ws) in zip(pipe.run(jobs), want):
		assert np.array_equal(res.dewarped, wd)
		assert len(res.strips) == len(ws)
		for a, b in zip(res.strips, ws):
			assert a.shape == b.shape and np.array_equal(a, b)
		n_seen += 1
	assert n_seen == len(jobs)
	# strips only (the OCR stage never touches the dewarped image): same strips, no page
	for res, (wd, ws) in zip(PagePipeline(line_height=48, binarize=spec, dewarped=False).run(jobs), want):
		assert res.dewarped is None and len(res.strips) == len(ws)
		for a, b in zip(res.strips, ws):
			assert a.shape == b.shape and np.array_equal(a, b)


def test_page_pipeline_with_column_descriptors(small):
	"""Table rows split into column cells travel through the pipeline as og_line descriptors."""
	from origami_b200.batch.core.pipeline import PagePipeline
	block = Block(small.pg, None, Stage.RELIABLE)
	recs, want = [], []
	for ln in small.lines[:12]:
		line = Line(block, ln["p"], ln["right"], ln["up"])
		x0 = ln["p"][0] + 0.3 * ln["right"][0]
		for col in ((None, x0), (x0, None)):
			recs.append(line.record(ysteps=48, column=col))
			want.append(reference_path.line_image(small.pil, small.ref, ln["p"], ln["right"], ln["up"], 48, column=col))
	pipe = PagePipeline(line_height=48)
	(res,) = list(pipe.run([(small.page, small.g, small.s, ops.pack_lines(recs))]))
	assert len(res.strips) == len(want)
	for a, b in zip(res.strips, want):
		assert a.shape == b.shape and np.array_equal(a, b)


def test_page_pipeline_reruns_a_page_whose_strip_estimate_fell_short(small, monkeypatch):
	"""The strip buffer is sized from a host-side estimate of the strip widths; when that is too
	small the kernels skip what does not fit (no write past the buffer) and the page runs again
	with the size the device computed."""
	from origami_b200.batch.core import pipeline
	lines = small.lines[:20]
	arr = np.array([[l["p"], l["right"], l["up"]] for l in lines], dtype=np.float64).reshape(-1, 3, 2)
	want = [reference_path.line_image(small.pil, small.ref, l["p"], l["right"], l["up"], 48, binarize=("otsu",)) for l in lines]
	good = pipeline._estimate_xsteps(small.g, small.s, ops.pack_line_arrays(arr[:, 0], arr[:, 1], arr[:, 2], ysteps=48))
	monkeypatch.setattr(pipeline, "_estimate_xsteps", lambda grid, cell, desc: np.full(desc.shape[0], 3, dtype=np.int64))
	pipe = pipeline.PagePipeline(line_height=48, binarize="otsu")
	results = list(pipe.run([(small.page, small.g, small.s, arr)] * 2))
	for res in results:
		assert len(res.strips) == len(want)
		for a, b in zip(res.strips, want):
			assert a.shape == b.shape and np.array_equal(a, b)
		# and the real estimate is tight: never below, at most a few columns above what the device derived
		slack = good - res.steps[:, 0]
		assert (slack >= 0).all() and (slack <= 16).all()


def test_page_pipeline_without_lines_and_without_dewarped_page(small):
	from origami_b200.batch.core.pipeline import PagePipeline
	empty = np.zeros((0, 3, 2), dtype=np.float64)
	for dewarped in (True, False):
		(res,) = list(PagePipeline(line_height=48, dewarped=dewarped).run([(small.page, small.g, small.s, empty)]))
		assert len(res.strips) == 0 and res.steps.shape == (0, 2)
		assert (res.dewarped is None) == (not dewarped)
		if dewarped:
			assert np.array_equal(res.dewarped, exact.remap_linear(small.page, exact.expand_map(small.g, small.s)))


def test_objects_on_a_second_device_while_the_first_is_current():
	"""ops.* run on the device that owns their tensors, whatever device is current (needs 2 GPUs)."""
	if torch.cuda.device_count() < 2:
		pytest.skip("needs two GPUs")
	W, H, s = 400, 300, 25
	page = synth.make_page(W, H, seed=2, kind="noise")
	g = synth.make_grid(W, H, s, seed=2)
	dense = exact.expand_map(g, s)
	torch.cuda.set_device(0)
	grid = Grid(g, s, device="cuda:1")
	pg = Page(page, grid)
	assert np.array_equal(np.array(pg.dewarped), exact.remap_linear(page, dense))
	assert np.array_equal(grid.points("full").view(np.uint32), dense.view(np.uint32))
	lines = synth.make_lines(dense.shape[1], dense.shape[0], 12, seed=2, dpi=150)
	block = Block(pg, None, Stage.RELIABLE)
	for ln in lines[:4]:
		im = Line(block, ln["p"], ln["right"], ln["up"]).image(48, binarizer=binarize.otsu())
		want = exact.binarize(exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48), ("otsu",))
		assert np.array_equal(np.array(im), want)
	assert torch.cuda.current_device() == 0
	with pytest.raises(ValueError):
		ops.remap_page(ops.to_device_image(page, torch.device("cuda:0")), grid.device_grid, s)
