"""GPU parity: line kernels (A5, A8, A9, A15, A21) through the C ABI against the oracle."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import ops, synth

pytestmark = pytest.mark.gpu


def _setup(cuda, width=600, height=800, cell=25, n_lines=80, seed=1, dpi=300, kind="noise"):
	page = synth.make_page(width, height, seed=seed, kind=kind)
	grid = synth.make_grid(width, height, cell, seed=seed)
	dense = exact.expand_map(grid, cell)
	lines = synth.make_lines(dense.shape[1], dense.shape[0], n_lines, seed=seed, dpi=dpi)
	return page, grid, dense, lines


def _records(lines, **kw):
	recs = []
	for ln in lines:
		r = dict(p=ln["p"], right=ln["right"], up=ln["up"])
		r.update(kw)
		recs.append(r)
	return recs


def test_inverse_yx_bit_exact(cuda):
	_, grid, dense, _ = _setup(cuda)
	rng = np.random.default_rng(0)
	Hd, Wd = dense.shape[:2]
	pts = np.stack([rng.uniform(-40, Hd + 40, 20000), rng.uniform(-40, Wd + 40, 20000)], axis=1)
	pts[:200] = np.rint(pts[:200])  # exactly integral coordinates
	want = exact.inverse_yx(dense, pts)
	got = ops.inverse_yx(torch.from_numpy(grid).to(cuda), 25, torch.from_numpy(pts).to(cuda)).cpu().numpy()
	assert np.array_equal(got, want)


@pytest.mark.parametrize("use_map", [False, True])
@pytest.mark.parametrize("target_height", [48, 56, 7])
def test_strips_bit_exact(cuda, target_height, use_map):
	page, grid, dense, lines = _setup(cuda)
	g = torch.from_numpy(grid).to(cuda)
	dmap = ops.expand_map(g, 25) if use_map else None
	src = ops.to_device_image(page, cuda)
	la = ops.pack_lines(_records(lines, ysteps=target_height))
	plan = ops.line_steps(g, 25, ops.lines_to_device(la, cuda), len(lines))
	steps = plan.steps.cpu().numpy()
	offs = plan.pix_offsets.cpu().numpy()
	out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	bbox = torch.zeros((len(lines), 4), dtype=torch.int32, device=cuda)
	ops.extract_lines(src, g, 25, plan, out, bbox, dense_map=dmap)
	out = out.cpu().numpy()
	bbox = bbox.cpu().numpy()
	for i, ln in enumerate(lines):
		wg = exact.warped_grid(dense, ln["p"], ln["right"], ln["up"], ysteps=target_height)
		want = exact.remap_linear(page, wg)[::-1]
		assert tuple(steps[i]) == (want.shape[1], want.shape[0]), i
		got = out[offs[i]:offs[i + 1]].reshape(want.shape)
		assert np.array_equal(got, want), i
		assert tuple(bbox[i]) == exact.grid_bbox(wg), i


@pytest.mark.parametrize("use_map", [False, True])
def test_warped_grids_half_res_bit_exact(cuda, use_map):
	_, grid, dense, lines = _setup(cuda, seed=4)
	g = torch.from_numpy(grid).to(cuda)
	dmap = ops.expand_map(g, 25) if use_map else None
	la = ops.pack_lines(_records(lines, xres=0.5, yres=0.5))
	plan = ops.line_steps(g, 25, ops.lines_to_device(la, cuda), len(lines))
	offs = plan.pix_offsets.cpu().numpy()
	steps = plan.steps.cpu().numpy()
	out = torch.zeros((int(offs[-1]) * 2,), dtype=torch.float32, device=cuda)
	ops.warped_grids(g, 25, plan, out, dense_map=dmap)
	out = out.cpu().numpy()
	for i, ln in enumerate(lines):
		want = exact.warped_grid(dense, ln["p"], ln["right"], ln["up"], xres=0.5, yres=0.5)
		assert tuple(steps[i]) == (want.shape[1], want.shape[0]), i
		got = out[2 * offs[i]:2 * offs[i + 1]].reshape(want.shape)
		assert np.array_equal(got, want), i


@pytest.mark.parametrize("use_map", [False, True])
def test_column_clipped_lines(cuda, use_map):
	"""Table rows split into column cells (LineRewriter + Line._position)."""
	page, grid, dense, lines = _setup(cuda, seed=6)
	g = torch.from_numpy(grid).to(cuda)
	dmap = ops.expand_map(g, 25) if use_map else None
	src = ops.to_device_image(page, cuda)
	recs, cases = [], []
	for ln in lines[:30]:
		x0 = ln["p"][0] + 0.2 * ln["right"][0]
		x1 = ln["p"][0] + 0.7 * ln["right"][0]
		for col in ((None, x0), (x0, x1), (x1, None)):
			cp, cright, _, xres = exact.line_position(ln["p"], ln["right"], ln["up"], 1, col)
			recs.append(dict(p=ln["p"], right=ln["right"], up=ln["up"], cp=cp, cright=cright, xres=xres, ysteps=48))
			cases.append((ln, col))
	la = ops.pack_lines(recs)
	plan = ops.line_steps(g, 25, ops.lines_to_device(la, cuda), len(recs))
	offs = plan.pix_offsets.cpu().numpy()
	out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, g, 25, plan, out, dense_map=dmap)
	out = out.cpu().numpy()
	for i, (ln, col) in enumerate(cases):
		want = exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48, column=col)
		assert offs[i + 1] - offs[i] == want.size, i
		assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), i


@pytest.mark.parametrize("use_map", [False, True])
def test_label_counts_bit_exact(cuda, use_map):
	page, grid, dense, lines = _setup(cuda, seed=8)
	labels = synth.make_labels(320, 600, seed=3)
	g = torch.from_numpy(grid).to(cuda)
	lab = ops.to_device_image(labels, cuda)
	la = ops.pack_lines(_records(lines, xres=0.5, yres=0.5))
	plan = ops.line_steps(g, 25, ops.lines_to_device(la, cuda), len(lines))
	dmap = ops.expand_map(g, 25) if use_map else None
	counts, overflow = ops.sample_labels(lab, page.shape, g, 25, plan, 4, dense_map=dmap)
	counts = counts.cpu().numpy()
	assert int(overflow.sum()) == 0
	for i, ln in enumerate(lines):
		want = exact.sample_label_counts(labels, page.shape, dense, ln["p"], ln["right"], ln["up"], 4)
		assert np.array_equal(counts[i], want[:4]), i


def test_empty_and_degenerate_lines(cuda):
	_, grid, dense, lines = _setup(cuda)
	g = torch.from_numpy(grid).to(cuda)
	la = ops.pack_lines([])
	plan = ops.line_steps(g, 25, torch.zeros((8,), dtype=torch.uint8, device=cuda), 0)
	assert plan.pix_offsets.cpu().tolist() == [0]
	# a zero-length line still yields the reference's minimum of 2 samples
	rec = dict(p=[100.0, 100.0], right=[0.0, 0.0], up=[0.0, -20.0], ysteps=48)
	la = ops.pack_lines([rec])
	plan = ops.line_steps(g, 25, ops.lines_to_device(la, cuda), 1)
	want = exact.warped_grid(dense, rec["p"], rec["right"], rec["up"], ysteps=48)
	assert plan.steps.cpu().numpy()[0].tolist() == [want.shape[1], want.shape[0]]
