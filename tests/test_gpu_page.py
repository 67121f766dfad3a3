"""GPU parity: page kernels (A2, A3, A4) through the C ABI against the oracle."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import ops, synth

pytestmark = pytest.mark.gpu


def _case(width, height, cell, seed, kind="noise"):
	page = synth.make_page(width, height, seed=seed, kind=kind)
	grid = synth.make_grid(width, height, cell, seed=seed)
	return page, grid


@pytest.mark.parametrize("width,height,cell", [(600, 800, 25), (333, 257, 8), (640, 480, 7), (96, 64, 3), (1200, 900, 25)])
def test_expand_map_bit_exact(cuda, width, height, cell):
	_, grid = _case(width, height, cell, seed=3)
	want = exact.expand_map(grid, cell)
	got = ops.expand_map(torch.from_numpy(grid).to(cuda), cell).cpu().numpy()
	assert got.shape == want.shape
	assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("width,height,cell,kind", [
	(600, 800, 25, "noise"), (333, 257, 8, "noise"), (640, 480, 7, "text"), (1201, 903, 25, "noise")])
def test_remap_page_bit_exact(cuda, width, height, cell, kind):
	page, grid = _case(width, height, cell, seed=5, kind=kind)
	dense = exact.expand_map(grid, cell)
	want = exact.remap_linear(page, dense)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	map_out = torch.empty(dense.shape, dtype=torch.float32, device=cuda)
	dst, _ = ops.remap_page(src, g, cell, map_out=map_out)
	assert np.array_equal(map_out.cpu().numpy().view(np.uint32), dense.view(np.uint32))
	assert np.array_equal(dst.cpu().numpy(), want)
	# image-only mode gives the same pixels
	dst2, _ = ops.remap_page(src, g, cell)
	assert torch.equal(dst, dst2)
	# ... and so does the shared-memory staged variant, with and without the map
	map3 = torch.empty_like(map_out)
	dst3, _ = ops.remap_page(src, g, cell, map_out=map3, interpolation=ops.OG_INTER_LINEAR | ops.OG_FLAG_STAGE_SMEM)
	assert torch.equal(dst, dst3) and torch.equal(map3, map_out)
	dst4, _ = ops.remap_page(src, g, cell, interpolation=ops.OG_INTER_LINEAR | ops.OG_FLAG_STAGE_SMEM)
	assert torch.equal(dst, dst4)


def test_remap_page_unaligned_buffers(cuda):
	"""Contiguous (unpitched) torch tensors with odd widths take the scalar store path."""
	page, grid = _case(331, 207, 5, seed=9)
	dense = exact.expand_map(grid, 5)
	want = exact.remap_linear(page, dense)
	g = torch.from_numpy(grid).to(cuda)
	src = torch.from_numpy(page).to(cuda)
	dst = torch.empty(dense.shape[:2], dtype=torch.uint8, device=cuda)
	ops.remap_page(src, g, 5, dst=dst)
	assert np.array_equal(dst.cpu().numpy(), want)


@pytest.mark.parametrize("interp", [exact.INTER_NEAREST, exact.INTER_LINEAR])
def test_dewarp_image_other_resolution(cuda, interp):
	"""Dewarper.dewarp_image: label-sized image through the page's grid (A4)."""
	width, height, cell = 600, 800, 25
	_, grid = _case(width, height, cell, seed=11)
	labels = synth.make_labels(320, 600, seed=2) * 60
	want = exact.dewarp_image(labels, (width, height), grid, cell, interp)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(labels, cuda)
	dst, _ = ops.remap_page(src, g, cell, scale=(320 / width, 600 / height), interpolation=interp)
	assert np.array_equal(dst.cpu().numpy(), want)
	if interp == exact.INTER_LINEAR:
		dst2, _ = ops.remap_page(src, g, cell, scale=(320 / width, 600 / height), interpolation=interp | ops.OG_FLAG_STAGE_SMEM)
		assert torch.equal(dst, dst2)


def test_page_argument_errors(cuda):
	g = torch.zeros((4, 4, 2), dtype=torch.float32, device=cuda)
	with pytest.raises(ValueError):
		ops.expand_map(g, 0)
	with pytest.raises(ValueError):
		ops.expand_map(torch.zeros((1400, 4, 2), dtype=torch.float32, device=cuda), 25)  # >= 32767 rows
