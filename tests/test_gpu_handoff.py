"""GPU parity of the strip hand-off: the OCR stage's >= 6 x 6 filter and
padded batch (origami/batch/detect/ocr.py:238-268) assembled on the device, against the oracle."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import ops, synth
from origami_b200.batch.core.lines import LineExtractor
from origami_b200.core.block import Block, Line, Stage
from origami_b200.core.dewarp import Grid
from origami_b200.core.page import Page

pytestmark = pytest.mark.gpu


def _page(seed=5, W=700, H=500, n=60):
	page = synth.make_page(W, H, seed=seed, kind="text", dpi=300)
	g = synth.make_grid(W, H, 25, seed=seed)
	dense = exact.expand_map(g, 25)
	lines = synth.make_lines(dense.shape[1], dense.shape[0], n, seed=seed, dpi=300, table_fraction=0.0)
	return page, g, dense, lines


@pytest.mark.parametrize("height,pad,w_align", [(48, 0, 4), (48, 255, 1), (5, 0, 4)])
def test_pack_strips_bit_exact(cuda, height, pad, w_align):
	page, g, dense, lines = _page()
	recs = [dict(p=l["p"], right=l["right"], up=l["up"], ysteps=height) for l in lines]
	for k in (3, 11, 12, 40):  # strips narrower than the minimum
		recs[k]["xsteps"] = 1 + k % 5
	gt = torch.from_numpy(g).to(cuda)
	src = ops.to_device_image(page, cuda)
	plan = ops.line_steps(gt, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	steps, offs = ops.plan_to_host(plan)
	gray = torch.empty((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, gt, 25, plan, gray)
	strips = [exact.remap_linear(page, exact.warped_grid(dense, r["p"], r["right"], r["up"], xsteps=r.get("xsteps"), ysteps=height))[::-1] for r in recs]
	w_max = (max(s.shape[1] for s in strips) + w_align - 1) // w_align * w_align + (1 if w_align == 1 else 0)
	want, want_w, want_k = exact.ocr_batch(strips, height, w_max, pad_value=pad)
	batch, widths, kept, n_kept = ops.pack_strips(gray, plan, height, w_max, pad_value=pad)
	nk = int(n_kept.item())
	assert nk == len(want_k) and (nk == 0 if height < 6 else 0 < nk < len(recs))
	assert np.array_equal(kept[:nk].cpu().numpy(), want_k) and np.array_equal(widths[:nk].cpu().numpy(), want_w)
	assert np.array_equal(batch[:nk].cpu().numpy(), want)


def test_extract_device_through_the_line_extractor(cuda):
	page, g, dense, lines = _page(seed=8)
	pg = Page(page, Grid(g, 25))
	block = Block(pg, None, Stage.RELIABLE)
	ldict = dict((("regions", "TEXT", str(i), "0"), Line(block, l["p"], l["right"], l["up"], confidence=l["confidence"])) for i, l in enumerate(lines))
	ex = LineExtractor(dict(columns={}), 48, dict(binarize="otsu", do_not_dewarp=False, do_not_deskew=False), min_confidence=0.5)
	paths, batch, widths, empty = ex.extract_device(ldict)
	host = ex.extract_arrays(ldict)
	assert [p for p, _ in host] == paths and empty == []
	assert batch.is_cuda and batch.shape[0] == len(paths) and batch.shape[1] == 48
	b = torch.utils.dlpack.from_dlpack(batch.__dlpack__()).cpu().numpy()  # exportable to the recogniser's framework
	for j, (_, strip) in enumerate(host):
		assert int(widths[j]) == strip.shape[1]
		assert np.array_equal(b[j, :, :strip.shape[1]], strip) and (b[j, :, strip.shape[1]:] == 0).all()


@pytest.mark.parametrize("spec", ["", "otsu"])
def test_pipeline_hands_the_strips_over_on_the_device(cuda, spec):
	from origami_b200.batch.core.pipeline import PagePipeline
	jobs, want = [], []
	for k, (W, H, n) in enumerate([(700, 500, 40), (333, 421, 25), (512, 512, 0)]):
		page = synth.make_page(W, H, seed=70 + k, kind="text", dpi=300)
		g = synth.make_grid(W, H, 25, seed=70 + k)
		dense = exact.expand_map(g, 25)
		lines = synth.make_lines(dense.shape[1], dense.shape[0], n, seed=70 + k, dpi=300, table_fraction=0.0)
		arr = np.array([[l["p"], l["right"], l["up"]] for l in lines], dtype=np.float64).reshape(-1, 3, 2)
		jobs.append((page, g, 25, arr))
		want.append([exact.binarize(exact.dewarped_line_image(page, dense, l["p"], l["right"], l["up"], 48), ("otsu",) if spec else None) for l in lines])
	pipe = PagePipeline(line_height=48, binarize=spec, dewarped=False, hand_off="device")
	for res, strips in zip(pipe.run(jobs), want):
		b = res.strips.batch
		assert b.is_cuda and b.shape[0] == len(strips) == len(res.strips.widths)
		if not strips:
			continue
		ref, ref_w, ref_k = exact.ocr_batch(strips, 48, b.shape[2])
		assert np.array_equal(res.strips.widths, ref_w) and np.array_equal(res.strips.line_index, ref_k)
		assert np.array_equal(b.cpu().numpy(), ref)


@pytest.mark.parametrize("n,offset", [(0, 0), (1, 0), (7, 0), (8, 0), (31, 0), (32, 0), (33, 0), (1000, 0), (100003, 0), (4099, 3), (65536, 16)])
def test_pack_bits_is_numpy_packbits(cuda, n, offset):
	rng = np.random.RandomState(n + offset)
	px = rng.randint(0, 3, size=n + offset).astype(np.uint8) * rng.randint(1, 256, size=n + offset).astype(np.uint8)  # 0 and arbitrary non-zero bytes
	src = torch.from_numpy(px).to(cuda)[offset:]
	got = ops.pack_bits(src.contiguous() if offset == 0 else src, n).cpu().numpy() if n else np.zeros(0, np.uint8)
	assert np.array_equal(got, np.packbits(px[offset:] != 0))


def test_pipeline_bits_hand_off_equals_host_strips(cuda):
	from origami_b200.batch.core.pipeline import PagePipeline
	page, g, dense, lines = _page(seed=9, n=80)
	arr = np.array([[l["p"], l["right"], l["up"]] for l in lines], dtype=np.float64)
	jobs = [(page, g, 25, arr)] * 3
	for spec in ("otsu", "sauvola(window_size=15)"):
		host = [[s.copy() for s in res.strips] for res in PagePipeline(48, spec, depth=2, dewarped=False).run(jobs)]
		bits = list(PagePipeline(48, spec, depth=2, dewarped=False, hand_off="bits").run(jobs))
		assert len(bits) == len(host) == 3
		for res, want in zip(bits, host):
			assert len(res.strips) == len(want) == len(lines)
			for a, b in zip(res.strips, want):
				assert a.dtype == np.uint8 and np.array_equal(a, b)
			assert np.array_equal(res.strips[-1], want[-1]) and np.array_equal(res.strips[5:7][1], want[6])
	with pytest.raises(ValueError):
		PagePipeline(48, "", hand_off="bits")
