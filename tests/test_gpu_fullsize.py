"""GPU parity at the sizes BASELINE.json names (configs[0], [1], [3], [4]); every comparison is
against the oracle (oracle/exact.py), not against another engine of this library."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import ops, synth

pytestmark = pytest.mark.gpu


def _strips(cuda, page, g, s, lines, height, binarize=None, use_map=False):
	gt = torch.from_numpy(g).to(cuda)
	dmap = ops.expand_map(gt, s) if use_map else None
	src = ops.to_device_image(page, cuda)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=height) for l in lines])
	plan = ops.line_steps(gt, s, ops.lines_to_device(la, cuda), len(lines))
	offs = plan.pix_offsets.cpu().numpy()
	out = torch.empty((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, gt, s, plan, out, binarize=binarize, dense_map=dmap)
	return out.cpu().numpy(), offs, plan


@pytest.mark.parametrize("kind,use_map", [("noise", True), ("text", True), ("noise", False), ("text", False)])
def test_all_strips_of_the_stress_page_against_the_oracle(cuda, kind, use_map):
	"""configs[1] + configs[3]: every one of the 2000 strips of a 4800x7200 page, gray and Otsu'd,
	against exact.dewarped_line_image / exact.binarize -- through the engines the bench times (dense map
	given: screened map-gather + Otsu pass) and through the line-owning engine (no map, Otsu fused)."""
	W, H, s = 4800, 7200, 25
	page = synth.make_page(W, H, seed=3, kind=kind)
	g = synth.make_grid(W, H, s, seed=3)
	dense = exact.expand_map(g, s)
	lines = synth.make_lines(dense.shape[1], dense.shape[0], 2000, seed=3, dpi=600)
	gray, offs, plan = _strips(cuda, page, g, s, lines, 48, use_map=use_map)
	fused, offs2, _ = _strips(cuda, page, g, s, lines, 48, binarize=("otsu",), use_map=use_map)
	assert np.array_equal(offs, offs2)
	two_pass = ops.binarize_strips(torch.from_numpy(gray).to(cuda), plan, ops.OG_BIN_OTSU).cpu().numpy()
	for i, l in enumerate(lines):
		want = exact.dewarped_line_image(page, dense, l["p"], l["right"], l["up"], 48)
		assert np.array_equal(gray[offs[i]:offs[i + 1]].reshape(want.shape), want), i
		wb = exact.binarize(want, ("otsu",))
		assert np.array_equal(fused[offs[i]:offs[i + 1]].reshape(want.shape), wb), i
		assert np.array_equal(two_pass[offs[i]:offs[i + 1]].reshape(want.shape), wb), i


def test_config0_page_2400x3600_with_1200_lines(cuda):
	"""configs[0]: the reference's CPU-runnable case, 300 dpi, strips of height 48 and 56."""
	W, H, s = 2400, 3600, 25
	page = synth.make_page(W, H, seed=5, kind="text", dpi=300)
	g = synth.make_grid(W, H, s, seed=5)
	dense = exact.expand_map(g, s)
	gt = torch.from_numpy(g).to(cuda)
	dmap = torch.empty(dense.shape, dtype=torch.float32, device=cuda)
	dst, _ = ops.remap_page(ops.to_device_image(page, cuda), gt, s, map_out=dmap)
	assert np.array_equal(dmap.cpu().numpy().view(np.uint32), dense.view(np.uint32))
	assert np.array_equal(dst.cpu().numpy(), exact.remap_linear(page, dense))
	lines = synth.make_lines(dense.shape[1], dense.shape[0], 1200, seed=5, dpi=300)
	for height, use_map in ((48, True), (56, False)):
		got, offs, _ = _strips(cuda, page, g, s, lines, height, use_map=use_map)
		for i, l in enumerate(lines):
			want = exact.dewarped_line_image(page, dense, l["p"], l["right"], l["up"], height)
			assert np.array_equal(got[offs[i]:offs[i + 1]].reshape(want.shape), want), (height, i)


@pytest.mark.parametrize("kind,window", [("text", 15), ("noise", 15), ("text", 31), ("noise", 31)])
def test_sauvola_on_the_full_page(cuda, kind, window):
	"""configs[4]: Page.binarized / Regions.binarized on a 4800x7200 page."""
	img = synth.make_page(4800, 7200, seed=window, kind=kind)
	want = exact.binarize(img, ("sauvola", window))
	got = ops.sauvola(ops.to_device_image(img, cuda), window).cpu().numpy()
	assert np.array_equal(got, want)


@pytest.mark.parametrize("M", [1, 5])
def test_argmax_on_the_legacy_label_map_size(cuda, M):
	"""configs[4]: (M, 2400, 1280, 4) softmax tiles -> labels and class masks."""
	probs = synth.make_probabilities(M, 1280, 2400, 4, seed=M)
	want = exact.argmax_labels(probs)
	labels, masks = ops.argmax_labels(torch.from_numpy(probs).to(cuda), [1, 2, 3])
	assert np.array_equal(labels.cpu().numpy(), want)
	assert np.array_equal(masks.cpu().numpy(), exact.class_masks(want, [1, 2, 3]))


def test_otsu_on_the_full_page(cuda):
	img = synth.make_page(4800, 7200, seed=9, kind="text")
	dst, t = ops.otsu(ops.to_device_image(img, cuda))
	assert np.array_equal(dst.cpu().numpy(), exact.binarize(img, ("otsu",)))
	assert int(t.item()) == exact.threshold_otsu(img)
