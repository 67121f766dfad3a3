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


def _check_lines(cuda, page, grid, dense, recs, use_map, cell=25):
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	dmap = ops.expand_map(g, cell) if use_map else None
	plan = ops.line_steps(g, cell, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	offs = plan.pix_offsets.cpu().numpy()
	out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, g, cell, plan, out, dense_map=dmap)
	out = out.cpu().numpy()
	for i, r in enumerate(recs):
		wg = exact.warped_grid(dense, r["p"], r["right"], r["up"], xsteps=r.get("xsteps"), ysteps=r.get("ysteps"),
			xres=r.get("xres", 1), yres=r.get("yres", 1))
		want = exact.remap_linear(page, wg)[::-1]
		assert offs[i + 1] - offs[i] == want.size, (i, want.shape)
		assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), i


@pytest.mark.parametrize("use_map", [False, True])
def test_unusual_lines_take_the_general_paths(cuda, use_map):
	"""Steep, very tall, very high (more rows than the row table holds), upsampled, tiny and
	partly off-page lines: the staged engine must fall back to its unstaged path where the node
	region does not fit, and both engines must still match the oracle bit for bit."""
	page, grid, dense, _ = _setup(cuda, seed=13)
	Hd, Wd = dense.shape[:2]
	recs = [
		dict(p=[60.0, 400.0], right=[180.0, 170.0], up=[14.0, -15.0], ysteps=48),            # 43 degrees
		dict(p=[300.0, 700.0], right=[120.0, -3.0], up=[2.0, -420.0], ysteps=48),            # 420 px tall
		dict(p=[100.0, 300.0], right=[90.0, 1.0], up=[0.5, -60.0], ysteps=200),              # ysteps > row table
		dict(p=[200.0, 200.0], right=[40.0, 0.0], up=[0.0, -12.0], xres=3.0, yres=3.0),      # upsampled, derived steps
		dict(p=[350.5, 120.25], right=[3.0, 0.0], up=[0.0, -2.0], ysteps=5),                 # tiny
		dict(p=[-30.0, 40.0], right=[200.0, 2.0], up=[1.0, -70.0], ysteps=48),               # leaves the page (extrapolation)
		dict(p=[Wd - 90.0, Hd - 5.0], right=[150.0, 0.0], up=[0.0, -40.0], ysteps=48),       # beyond the right / bottom edge
		dict(p=[120.0, 500.0], right=[-100.0, 4.0], up=[1.0, 30.0], ysteps=33),              # reversed directions
		dict(p=[64.0, 256.0], right=[128.0, 0.0], up=[0.0, -32.0], ysteps=48, xsteps=300),   # fixed xsteps, integral lattice
	]
	_check_lines(cuda, page, grid, dense, recs, use_map)


def test_capacity_guard_skips_lines_that_do_not_fit(cuda):
	page, grid, dense, lines = _setup(cuda)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = _records(lines[:8], ysteps=48)
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	offs = plan.pix_offsets.cpu().numpy()
	cap = int(offs[5])  # room for the first five strips only
	for dmap in (None, ops.expand_map(g, 25)):
		out = torch.full((int(offs[-1]),), 7, dtype=torch.uint8, device=cuda)
		ops.extract_lines(src, g, 25, plan, out[:cap], dense_map=dmap)
		out = out.cpu().numpy()
		assert (out[cap:] == 7).all()  # nothing written past the capacity
		want = exact.dewarped_line_image(page, dense, lines[4]["p"], lines[4]["right"], lines[4]["up"], 48)
		assert np.array_equal(out[offs[4]:offs[5]].reshape(want.shape), want)


def test_degenerate_coordinates_are_flagged(cuda):
	_, grid, dense, _ = _setup(cuda)
	g = torch.from_numpy(grid).to(cuda)
	recs = [dict(p=[float("nan"), 10.0], right=[50.0, 0.0], up=[0.0, -20.0]),
		dict(p=[10.0, 10.0], right=[50.0, 0.0], up=[0.0, -20.0])]
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), 2)
	steps = plan.steps.cpu().numpy()
	assert steps[0].tolist() == [0, 0] and steps[1].min() >= 2


# ---------------------------------------------------------------------------------------
# The float32-screened strip path (dense map given, no bounding boxes asked for)
# ---------------------------------------------------------------------------------------
def _adversarial_maps(dense, seed=5):
	"""Dense maps that are NOT the expansion of a grid: rough ones (large node differences, so a
	wide screening interval), ones that leave the page, and ones whose values sit exactly on the
	1/64-px lattice where cv2.remap's rounding flips."""
	rng = np.random.default_rng(seed)
	f32 = np.float32
	return {
		"noise 0.5 px": (dense + rng.normal(0, 0.5, dense.shape)).astype(f32),
		"noise 30 px": (dense + rng.normal(0, 30, dense.shape)).astype(f32),
		"partly outside": (dense * f32(1.7)).astype(f32),
		"negative": (dense - f32(40)).astype(f32),
		"constant": np.full_like(dense, 123.5),
		"tiny": (dense * f32(1e-3)).astype(f32),
		"on the 1/64 lattice": (np.round(dense * 64) / 64).astype(f32),
		"on the rounding ties": (np.round(dense * 32) / 32 + f32(1 / 64)).astype(f32),
		"far outside": (dense * f32(1e6)).astype(f32),
		"across a float32 ulp step": (f32(256) + (dense - dense.mean()) * f32(0.01)).astype(f32),
		"steep": ((dense - dense.mean()) * f32(50) + f32(300)).astype(f32),
		"nearly flat": (f32(100) + dense * f32(1e-6)).astype(f32),
		"next to the ties": (np.round(dense * 32) / 32 + f32(1 / 64) + rng.normal(0, 1e-6, dense.shape)).astype(f32),
	}


@pytest.mark.parametrize("ysteps", [48, 7])
def test_screened_strips_on_adversarial_maps(cuda, ysteps):
	page, grid, dense, lines = _setup(cuda, n_lines=40, seed=3)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = _records(lines, ysteps=ysteps, xsteps=150)
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	offs = plan.pix_offsets.cpu().numpy()
	for name, m in _adversarial_maps(dense).items():
		out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
		ops.extract_lines(src, g, 25, plan, out, dense_map=torch.from_numpy(m).to(cuda), map_gather=True)
		out = out.cpu().numpy()
		for i, r in enumerate(recs):
			wg = exact.warped_grid(m, r["p"], r["right"], r["up"], xsteps=150, ysteps=ysteps)
			want = exact.remap_linear(page, wg)[::-1]
			assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), (name, i)


def test_screened_strips_equal_the_float64_path(cuda):
	"""A page-sized case: the screened kernel against the kernel that evaluates every sample in
	float64 (the bounding-box variant), on the grid's own map and on rough ones."""
	page, grid, dense, lines = _setup(cuda, width=1500, height=2000, n_lines=400, seed=21, dpi=600, kind="text")
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = _records(lines, ysteps=48)
	recs += [dict(p=[40.0, 900.0], right=[700.0, 650.0], up=[30.0, -32.0], ysteps=48),      # steep
		dict(p=[-50.0, 300.0], right=[400.0, 3.0], up=[1.0, -60.0], ysteps=51),               # off-page, odd height
		dict(p=[200.0, 1500.0], right=[900.0, -8.0], up=[2.0, -70.0], ysteps=300)]            # taller than the row table
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	total = int(plan.pix_offsets[-1].item())
	bbox = torch.zeros((len(recs), 4), dtype=torch.int32, device=cuda)
	maps = {"grid": ops.expand_map(g, 25)}
	maps.update({k: torch.from_numpy(v).to(cuda) for k, v in _adversarial_maps(dense).items()})
	for name, m in maps.items():
		a = torch.zeros((total,), dtype=torch.uint8, device=cuda)
		b = torch.zeros((total,), dtype=torch.uint8, device=cuda)
		ops.extract_lines(src, g, 25, plan, a, dense_map=m, map_gather=True)
		ops.extract_lines(src, g, 25, plan, b, bbox, dense_map=m)
		assert torch.equal(a, b), name
		# the work ticket behind the prefix sums is back at zero, so the plan can be used again
		assert int(plan.chunk_offsets[len(recs) + 1].item()) == 0
		a2 = torch.zeros_like(a)
		ops.extract_lines(src, g, 25, plan, a2, dense_map=m, map_gather=True)
		assert torch.equal(a, a2), name
		if name == "grid":  # the line-owning engine recomputes the nodes from the sparse grid: same bytes
			c = torch.zeros_like(a)
			ops.extract_lines(src, g, 25, plan, c)
			assert torch.equal(a, c)
			assert int(plan.chunk_offsets[len(recs) + 1].item()) == 0


def test_screened_strips_with_absurd_line_coordinates(cuda):
	"""Coordinates beyond 2^27 switch the whole line to the float64 path (the fixed-point trick of
	the screening needs |x| < 2^27); the result is whatever the reference's arithmetic gives."""
	page, grid, dense, _ = _setup(cuda, seed=4)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = [dict(p=[1.0e9, 50.0], right=[100.0, 0.0], up=[0.0, -20.0], ysteps=8, xsteps=64),
		dict(p=[100.0, 200.0], right=[3.0e8, 0.0], up=[0.0, -20.0], ysteps=8, xsteps=64),
		dict(p=[100.0, 200.0], right=[200.0, 0.0], up=[0.0, -20.0], ysteps=8, xsteps=64)]
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	total = int(plan.pix_offsets[-1].item())
	bbox = torch.zeros((len(recs), 4), dtype=torch.int32, device=cuda)
	m = ops.expand_map(g, 25)
	a = torch.zeros((total,), dtype=torch.uint8, device=cuda)
	b = torch.zeros((total,), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, g, 25, plan, a, dense_map=m, map_gather=True)
	ops.extract_lines(src, g, 25, plan, b, bbox, dense_map=m)
	assert total == 3 * 8 * 64 and torch.equal(a, b)
	c = torch.zeros_like(a)
	ops.extract_lines(src, g, 25, plan, c)  # the line-owning engine sends such lines through its float64 path
	assert torch.equal(a, c)
	want = exact.remap_linear(page, exact.warped_grid(dense, recs[2]["p"], recs[2]["right"], recs[2]["up"], xsteps=64, ysteps=8))[::-1]
	assert np.array_equal(a.cpu().numpy()[2 * 512:].reshape(want.shape), want)


def test_line_plan_large_batches_take_the_two_kernel_path(cuda):
	"""Up to 8192 lines the step counts and prefix sums come from one cluster launch, beyond that
	from two kernels: same numbers either way."""
	_, grid, dense, lines = _setup(cuda, n_lines=90, seed=8)
	g = torch.from_numpy(grid).to(cuda)
	recs = _records(lines)
	small = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	reps = 8192 // len(recs) + 1
	big = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs * reps), cuda), len(recs) * reps)
	s_small, s_big = small.steps.cpu().numpy(), big.steps.cpu().numpy()
	assert len(recs) * reps > 8192 and np.array_equal(s_big, np.tile(s_small, (reps, 1)))
	px = s_big[:, 0].astype(np.int64) * s_big[:, 1]
	assert np.array_equal(big.pix_offsets.cpu().numpy(), np.concatenate([[0], np.cumsum(px)]))
	chunks = (s_big[:, 0] + 63) // 64
	assert np.array_equal(big.chunk_offsets.cpu().numpy()[:-1], np.concatenate([[0], np.cumsum(chunks)]))
	assert int(big.chunk_offsets[-1].item()) == 0 and int(small.chunk_offsets[-1].item()) == 0
	for i, r in enumerate(recs):
		assert tuple(s_small[i]) == exact.line_steps(dense, r["p"], r["right"], r["up"])


def test_screened_strips_with_more_lines_than_the_offset_cache(cuda):
	"""More than 4096 lines: the prefix sums no longer fit the kernel's shared-memory cache and are
	searched in global memory; few pixels per line, so most tickets are row quarters."""
	page, grid, dense, lines = _setup(cuda, n_lines=60, seed=17)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = (_records(lines, ysteps=9, xsteps=40) + _records(lines, ysteps=48)) * 40  # 4800 lines
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	total = int(plan.pix_offsets[-1].item())
	m = ops.expand_map(g, 25)
	a = torch.zeros((total,), dtype=torch.uint8, device=cuda)
	b = torch.zeros((total,), dtype=torch.uint8, device=cuda)
	bbox = torch.zeros((len(recs), 4), dtype=torch.int32, device=cuda)
	ops.extract_lines(src, g, 25, plan, a, dense_map=m, map_gather=True)
	ops.extract_lines(src, g, 25, plan, b, bbox, dense_map=m)
	assert len(recs) > 4096 and torch.equal(a, b)
	c = torch.zeros_like(a)
	ops.extract_lines(src, g, 25, plan, c)
	assert torch.equal(a, c)
	offs = plan.pix_offsets.cpu().numpy()
	out = a.cpu().numpy()
	for i in (0, 59, 60, 119, len(recs) - 1):
		r = recs[i]
		want = exact.remap_linear(page, exact.warped_grid(dense, r["p"], r["right"], r["up"], xsteps=r.get("xsteps"), ysteps=r["ysteps"]))[::-1]
		assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), i


def test_screened_label_counts_on_adversarial_maps(cuda):
	"""Label sampling through the screened engine: custom dense maps (rough, off-page, on the rounding
	ties of the NEAREST lookup) against the oracle, overflow labels included; the ticket is reset."""
	page, grid, dense, lines = _setup(cuda, n_lines=40, seed=6)
	labels = synth.make_labels(320, 600, seed=4)
	labels[::7, ::5] = 9  # labels beyond n_classes -> overflow
	g = torch.from_numpy(grid).to(cuda)
	lab = ops.to_device_image(labels, cuda)
	recs = _records(lines, xsteps=70, ysteps=13)
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	maps = _adversarial_maps(dense)
	sx, sy = np.float32(labels.shape[1] / page.shape[1]), np.float32(labels.shape[0] / page.shape[0])
	maps["on the nearest ties"] = np.dstack([(np.round(dense[..., 0] * sx) + 0.5) / sx, (np.round(dense[..., 1] * sy) + 0.5) / sy]).astype(np.float32)
	maps["grid"] = dense
	for name, m in maps.items():
		counts, overflow = ops.sample_labels(lab, page.shape, g, 25, plan, 4, dense_map=torch.from_numpy(np.ascontiguousarray(m)).to(cuda))
		counts, overflow = counts.cpu().numpy(), overflow.cpu().numpy()
		assert int(plan.chunk_offsets[len(recs) + 1].item()) == 0
		for i, r in enumerate(recs):
			wg = exact.warped_grid(m, r["p"], r["right"], r["up"], xsteps=70, ysteps=13)
			got = exact.remap_nearest(labels, exact.scale_map(wg, page.shape[1], page.shape[0], labels.shape[1], labels.shape[0]))
			want = np.bincount(got.reshape(-1), minlength=10)
			assert np.array_equal(counts[i], want[:4]) and overflow[i] == want[4:].sum(), (name, i)


# ---------------------------------------------------------------------------------------
# The line-owning engine (og_strips.cu): nodes recomputed from the sparse grid, float32 screening
# with one error radius per chunk, Otsu fused
# ---------------------------------------------------------------------------------------
def _adversarial_grids(grid, seed=11):
	"""Sparse grids that stress the screening: rough ones (large node differences, so a wide
	interval and a large staged rectangle), ones that leave the page or the float32 range where
	cv2.remap's fixed point is defined, flat ones (every sample in the same source pixel), and
	ones whose dense nodes sit on cv2.remap's rounding ties."""
	rng = np.random.default_rng(seed)
	f32 = np.float32
	mean = grid.mean(axis=(0, 1), keepdims=True)
	return {
		"grid": grid,
		"noise 2 px": (grid + rng.normal(0, 2, grid.shape)).astype(f32),
		"noise 40 px": (grid + rng.normal(0, 40, grid.shape)).astype(f32),
		"partly outside": (grid * f32(1.7)).astype(f32),
		"negative": (grid - f32(60)).astype(f32),
		"constant": np.full_like(grid, 123.5),
		"tiny": (grid * f32(1e-3)).astype(f32),
		"integral nodes": np.round(grid).astype(f32),
		"on the rounding ties": (np.round(grid) + f32(1 / 64)).astype(f32),
		"far outside": (grid * f32(1e6)).astype(f32),
		"across a float32 ulp step": (f32(256) + (grid - mean) * f32(0.01)).astype(f32),
		"steep": ((grid - mean) * f32(20) + f32(300)).astype(f32),
		"nearly flat": (f32(100) + grid * f32(1e-6)).astype(f32),
		"not finite": np.where(rng.random(grid.shape) < 0.02, f32(np.inf), grid).astype(f32),
	}


@pytest.mark.parametrize("ysteps,binarize", [(48, None), (48, ("otsu",)), (7, None), (31, ("sauvola", 9))])
def test_line_owning_engine_on_adversarial_grids(cuda, ysteps, binarize):
	page, grid, _, lines = _setup(cuda, n_lines=40, seed=3, kind="text")
	src = ops.to_device_image(page, cuda)
	recs = _records(lines, ysteps=ysteps, xsteps=150)
	recs.append(dict(p=[60.0, 400.0], right=[180.0, 170.0], up=[14.0, -15.0], ysteps=ysteps, xsteps=150))  # 43 degrees
	la = ops.lines_to_device(ops.pack_lines(recs), cuda)
	for name, gr in _adversarial_grids(grid).items():
		g = torch.from_numpy(gr).to(cuda)
		with np.errstate(invalid="ignore", over="ignore"):
			dense = exact.expand_map(gr, 25)
		plan = ops.line_steps(g, 25, la, len(recs))
		offs = plan.pix_offsets.cpu().numpy()
		out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
		thr = torch.full((len(recs),), -1, dtype=torch.int32, device=cuda)
		ops.extract_lines(src, g, 25, plan, out, binarize=binarize, thresholds=thr)
		out = out.cpu().numpy()
		with np.errstate(invalid="ignore", over="ignore"):
			for i, r in enumerate(recs):
				wg = exact.warped_grid(dense, r["p"], r["right"], r["up"], xsteps=150, ysteps=ysteps)
				want = exact.binarize(exact.remap_linear(page, np.nan_to_num(wg, nan=-1e9, posinf=1e9, neginf=-1e9))[::-1], binarize)
				assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), (name, i)
		if binarize == ("otsu",):
			t = thr.cpu().numpy()
			assert ((t >= 0) & (t <= 255)).all()


def test_line_owning_engine_keeps_gray_beside_otsu(cuda):
	page, grid, dense, lines = _setup(cuda, n_lines=30, seed=9, kind="text")
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	recs = _records(lines, ysteps=48)
	plan = ops.line_steps(g, 25, ops.lines_to_device(ops.pack_lines(recs), cuda), len(recs))
	offs = plan.pix_offsets.cpu().numpy()
	binar = torch.zeros((int(offs[-1]) + 3,), dtype=torch.uint8, device=cuda)[3:]  # odd alignment of the packed buffer
	gray = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	thr = torch.zeros((len(recs),), dtype=torch.int32, device=cuda)
	ops.extract_lines(src, g, 25, plan, binar, binarize=("otsu",), gray_out=gray, thresholds=thr)
	binar, gray, thr = binar.cpu().numpy(), gray.cpu().numpy(), thr.cpu().numpy()
	for i, ln in enumerate(lines):
		want = exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48)
		assert np.array_equal(gray[offs[i]:offs[i + 1]].reshape(want.shape), want), i
		assert np.array_equal(binar[offs[i]:offs[i + 1]].reshape(want.shape), exact.binarize(want, ("otsu",))), i
		try:
			tw = exact.threshold_otsu(want)
		except ValueError:
			tw = 128
		assert thr[i] == tw
