"""CPU tests of the host-side mirror of the reference interface (no GPU needed)."""

import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from oracle import exact
from origami_b200 import shard, synth
from origami_b200.batch.core.lines import LineExtractor, LineRewriter
from origami_b200.core import binarize
from origami_b200.core.block import Block, Line, Stage, affine_transform, position
from origami_b200.core.dewarp import Grid

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_stage_is_dewarped():
	assert not Stage.WARPED.is_dewarped and Stage.DEWARPED.is_dewarped and Stage.RELIABLE.is_dewarped


def test_grid_container_and_zip_roundtrip(tmp_path):
	g = synth.make_grid(300, 200, 25, seed=2)
	grid = Grid(g, 25)
	assert grid.geometry.size == (g.shape[1] * 25, g.shape[0] * 25)
	assert grid.resolution == 25
	assert grid.points("sample") is not None and np.array_equal(grid.points("sample"), g)
	path = tmp_path / "dewarp.zip"
	grid.save(path)
	back = Grid.open(path)
	assert np.array_equal(back.points("sample"), g) and back.resolution == 25
	with pytest.raises(ValueError):
		grid.points("half")
	with pytest.raises(ValueError):
		Grid(np.zeros((4, 4)), 25)


def test_binarizer_spec_strings():
	assert binarize.from_string("otsu").spec == ("otsu",)
	assert binarize.from_string("otsu()").spec == ("otsu",)
	assert binarize.from_string("sauvola").spec == ("sauvola", 15)
	assert binarize.from_string("sauvola(window_size=31)").spec == ("sauvola", 31)
	with pytest.raises(ValueError):
		binarize.from_string("niblack(3)")


def test_position_matches_the_independent_oracle():
	"""Line._position: the product evaluates the hit parametrically in exact rational arithmetic,
	the oracle with GEOS's orientation tests + determinant form in 60-digit decimals; they must agree
	to the last bit, also when a column boundary passes through an end point of the baseline, misses
	it, or the baseline runs right to left."""
	rng = np.random.default_rng(0)
	cases = 0
	for k in range(120):
		p = rng.uniform(10, 500, 2)
		right = np.array([rng.uniform(50, 400), rng.uniform(-3, 3)])
		if k % 7 == 0:
			right[0] = -right[0]  # right to left
		if k % 11 == 0:
			right[1] = 0.0
		up = np.array([rng.uniform(-1, 1), -rng.uniform(10, 40)])
		lo, hi = sorted((p[0], p[0] + right[0]))
		x0 = lo + rng.uniform(0.05, 0.45) * (hi - lo)
		x1 = lo + rng.uniform(0.55, 0.95) * (hi - lo)
		for col in (None, (None, x0), (x0, x1), (x1, None), (lo - 50, x1), (lo, hi), (x0, hi + 3), (float(np.float32(x0)), None)):
			a = position(p, right, up, 1, col)
			b = exact.line_position(p, right, up, 1, col)
			for u, v in zip(a, b):
				assert np.array_equal(np.asarray(u), np.asarray(v)), (k, col)
			cases += 1
	assert cases == 960
	# a boundary through an end point returns that end point's own coordinates
	p, right, up = np.array([10.25, 20.1]), np.array([100.3, 0.7]), np.array([0.0, -10.0])
	q0, r, _, _ = position(p, right, up, 1, (None, 60.0))
	assert np.array_equal(q0, p) and r[0] == 60.0 - p[0]
	q0, r, _, _ = position(p, right, up, 1, (50.0, None))
	assert np.array_equal(q0 + r, p + right) or abs((q0 + r - (p + right))[1]) < 1e-13


def test_line_rewriter_splits_table_rows():
	tables = dict(columns={"regions/TABULAR/3.1.2.1": [120.0, 260.0]})
	rw = LineRewriter(tables)
	lines = {("regions", "TEXT", "7", "0"): "a", ("regions", "TABULAR", "3.1.2.1", "4"): "b"}
	parts = rw(lines)
	assert parts[0] == (("regions", "TEXT", "7", "0"), "a", None)
	cols = parts[1:]
	assert [c[2] for c in cols] == [(None, 120.0), (120.0, 260.0), (260.0, None)]
	assert [c[0] for c in cols] == [("regions", "TABULAR", "3.1.5.%d" % k, "0") for k in (1, 2, 3)]
	with pytest.raises(RuntimeError):
		LineRewriter(dict(columns={"regions/TABULAR/3": [1.0]}))({("regions", "TABULAR", "3", "0"): "x"})


def test_line_extractor_filters_confidence():
	ex = LineExtractor(dict(columns={}), 48, dict(binarize="", do_not_dewarp=False, do_not_deskew=False), min_confidence=0.5)
	block = Block(None, None, Stage.RELIABLE)
	lines = {
		("regions", "TEXT", "1", "0"): Line(block, [0, 10], [100, 0], [0, -10], confidence=0.9),
		("regions", "TEXT", "1", "1"): Line(block, [0, 30], [100, 0], [0, -10], confidence=0.2),
		("illustrations", "X", "2", "0"): Line(block, [0, 50], [100, 0], [0, -10], confidence=1.0),
	}
	parts = ex._filtered_parts(lines, ignored=lambda k: k[0] == "illustrations")
	assert [p[0] for p in parts] == [("regions", "TEXT", "1", "0")]
	with pytest.raises(ValueError):
		LineExtractor(dict(columns={}), 48, dict(binarize="foo"))


def test_line_record_and_info():
	block = Block(None, None, Stage.RELIABLE)
	ln = Line(block, [10.0, 20.0], [200.0, 1.0], [0.1, -30.0], tesseract_data=dict(baseline=[[0, 0], [1, 1]], descent=1, ascent=2, height=3))
	rec = ln.record(ysteps=48, column=(50.0, 150.0))
	assert rec["ysteps"] == 48 and rec["xsteps"] is None
	assert rec["cp"][0] == 50.0 and abs(rec["cright"][0] - 100.0) < 1e-12
	assert abs(rec["xres"] - 0.5) < 1e-12
	info = ln.info
	assert info["p"] == [10.0, 20.0] and info["tesseract_data"]["height"] == 3


def test_affine_transform_is_cv2():
	cv2 = pytest.importorskip("cv2")
	src = np.array([[10.5, 20.25], [210.0, 24.0], [9.0, -8.0]], dtype=np.float32)
	dst = np.array([[0, 47], [200, 47], [0, 0]], dtype=np.float32)
	assert np.allclose(affine_transform(src, dst), cv2.getAffineTransform(src, dst), atol=1e-9)


def test_pages_of_rank_partition():
	for n, world in [(256, 8), (7, 2), (3, 4), (0, 2)]:
		seen = sorted(p for r in range(world) for p in shard.pages_of_rank(n, r, world))
		assert seen == list(range(n))
	assert shard.pages_of_rank(256, 3, 8)[:3] == [3, 11, 19]
	assert shard.max_over_ranks(1.5) == 1.5  # single process: no group


def test_sharding_two_ranks_gloo():
	"""world_size 2 on CPU: page shards are disjoint and the job time is the max over ranks."""
	with tempfile.TemporaryDirectory() as tmp:
		cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
			"--master-addr", "127.0.0.1", "--master-port", "29617",
			os.path.join(ROOT, "tests", "_gloo_worker.py"), tmp]
		res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
		assert res.returncode == 0, res.stderr[-2000:]
		outs = [json.load(open(os.path.join(tmp, "rank%d.json" % r))) for r in range(2)]
	assert outs[0]["pages"] == [0, 2, 4, 6] and outs[1]["pages"] == [1, 3, 5]
	assert all(o["tmax"] == 11.0 and o["total"] == 7.0 and o["world"] == 2 for o in outs)


def test_bench_reference_arm_noop_on_other_ranks():
	env = dict(os.environ, RANK="1", WORLD_SIZE="2")
	res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
		capture_output=True, text=True, env=env, timeout=120)
	assert res.returncode == 0 and res.stdout.strip() == ""


def test_pipeline_strip_width_estimate_bounds_the_device_value():
	"""PagePipeline sizes its strip buffer from a host-side estimate of the strip widths (float64
	bilinear of the sparse grid): never below what og_line_steps derives (the oracle's value), and
	at most four columns above it, also for lines at the border and lines that leave the page."""
	from origami_b200 import ops
	from origami_b200.batch.core import pipeline
	grid = synth.make_grid(900, 1200, 25, seed=5)
	dense = exact.expand_map(grid, 25)
	lines = synth.make_lines(dense.shape[1], dense.shape[0], 120, seed=5, dpi=300)
	lines += [dict(p=[-40.0, 300.0], right=[400.0, 5.0], up=[1.0, -30.0]),
		dict(p=[dense.shape[1] - 100.0, 200.0], right=[300.0, -4.0], up=[0.0, -25.0])]
	desc = ops.pack_line_arrays([l["p"] for l in lines], [l["right"] for l in lines], [l["up"] for l in lines], ysteps=48)
	est = pipeline._estimate_xsteps(grid, 25, desc)
	want = np.array([exact.line_steps(dense, l["p"], l["right"], l["up"], ysteps=48)[0] for l in lines])
	assert (est >= want).all() and (est - want <= 4).all(), ((est - want).min(), (est - want).max())
	desc["xsteps"][:10] = 77  # fixed widths are taken as they are
	assert (pipeline._estimate_xsteps(grid, 25, desc)[:10] == 77).all()


def test_lingrid_extrapolation_points_match_the_oracle():
	"""The hull / box / corner points of lingrid's extrapolating interpolator (host geometry):
	product construction against the oracle's restatement, on a warped lattice and on a small cloud."""
	from origami_b200.core.lingrid import extrapolation_points
	g = synth.make_grid(600, 400, 25, seed=3).astype(np.float64)
	src = g.reshape(-1, 2)
	h, w = g.shape[:2]
	tgt = np.flip(np.dstack(np.mgrid[0:h, 0:w]), axis=-1).reshape(-1, 2) * 25.0
	bounds = (src[:, 0].min(), src[:, 1].min(), src[:, 0].max(), src[:, 1].max())
	rng = np.random.default_rng(5)
	cloud = rng.uniform(10, 90, (12, 2))
	for pts, vals, b in ((src, tgt, bounds), (cloud, rng.uniform(-3, 3, (12, 2)), (0, 0, 100, 100))):
		ep, ev, collinear = extrapolation_points(pts, vals, b)
		wp, wv = exact.lininterp_points(pts, vals, b)
		assert not collinear and ep.shape == wp.shape
		assert np.allclose(ep, wp, rtol=0, atol=1e-9) and np.allclose(ev, wv, rtol=0, atol=1e-9)


def test_segment_zip_roundtrip_and_line_png_export(tmp_path):
	"""segment.zip (paletted PNG + class table, core/segment.py:45-80) and the PNG export of strips
	(batch/export/lines.py:57-67): written and read back byte for byte."""
	import io
	import zipfile
	import PIL.Image
	from origami_b200.core.segment import Prediction, PredictorType, Segmentation, write_line_images
	labels = synth.make_labels(128, 240, seed=1)
	seps = (synth.make_labels(128, 240, seed=2) == 2).astype(np.uint8)
	preds = [Prediction(PredictorType.REGION, "regions", labels, PredictorType.REGION.classes(dict(BACKGROUND=0, TEXT=1, TABULAR=2, ILLUSTRATION=3))),
		Prediction(PredictorType.SEPARATOR, "separators", seps, PredictorType.SEPARATOR.classes(dict(BACKGROUND=0, H=1)))]
	path = tmp_path / "segment.zip"
	Segmentation(preds).save(path)
	back = Segmentation.open(path)
	assert back.size == (128, 240) and [p.name for p in back.predictions] == ["regions", "separators"]
	for a, b in zip(preds, back.predictions):
		assert np.array_equal(a.labels, b.labels) and b.labels.dtype == np.uint8
		assert [(m.name, m.value) for m in a.classes] == [(m.name, m.value) for m in b.classes]
		assert b.background_label.value == 0 and b.type == a.type
	strips = [(("regions", "TEXT", "1", str(k)), (np.arange(48 * (20 + k)) % 251).astype(np.uint8).reshape(48, 20 + k)) for k in range(3)]
	zpath = tmp_path / "lines.zip"
	with zipfile.ZipFile(zpath, "w") as zf:
		write_line_images(zf, strips)
	with zipfile.ZipFile(zpath) as zf:
		assert zf.namelist() == ["regions/TEXT/1/%d.png" % k for k in range(3)]
		for (stem, arr), name in zip(strips, zf.namelist()):
			assert np.array_equal(np.array(PIL.Image.open(io.BytesIO(zf.read(name)))), arr)


def test_flow_samples_zip_roundtrip(tmp_path):
	"""flow.zip samples (origami/core/dewarp.py:43-74)."""
	import zipfile
	from origami_b200.core.gridfactory import Samples
	from origami_b200.core.math import Geometry
	s = Samples(Geometry(2400, 3600))
	s.extend([np.array([10.0, 20.0]), np.array([300.5, 40.25])], [0.01, -0.02])
	s.append(np.array([5.0, 6.0]), 0.5)
	path = tmp_path / "flow.zip"
	with zipfile.ZipFile(path, "w") as zf:
		s.save(zf, "h")
		Samples(Geometry(2400, 3600)).save(zf, "v")
	with zipfile.ZipFile(path) as zf:
		h, v = Samples.open(zf, "h"), Samples.open(zf, "v")
	assert len(h) == 3 and np.allclose(np.array(h.points), np.array(s.points)) and np.allclose(h.values, s.values)
	assert h.geometry.size == (2400, 3600) and len(v.values) == 0 and h.std == 0


def test_packed_strip_list_unpacks_what_strip_list_views():
	from origami_b200.core.block import PackedStripList, StripList
	rng = np.random.RandomState(0)
	steps = np.array([[5, 3], [7, 1], [1, 1], [13, 4], [8, 8], [3, 5]], np.int32)
	offs = np.concatenate([[0], np.cumsum(steps[:, 0] * steps[:, 1])]).astype(np.int64)
	flat = (rng.randint(0, 2, size=int(offs[-1])) * 255).astype(np.uint8)
	a, b = StripList(flat, offs, steps), PackedStripList(np.packbits(flat != 0), offs, steps)
	assert len(a) == len(b) == 6
	assert all(x.shape == y.shape and np.array_equal(x, y) for x, y in zip(a, b))
	assert np.array_equal(a[3], b[3]) and np.array_equal(a[-1], b[-1]) and np.array_equal(a[1:3][1], b[1:3][1])
