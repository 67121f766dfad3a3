"""GPU parity: binarizers (A12), label argmax / masks (A18, A19), affine deskew (A11)."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import ops, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,window", [((300, 420), 15), ((97, 131), 3), ((64, 64), 31), ((40, 500), 15), ((5, 9), 7),
	((200, 700), 31), ((150, 600), 41)])
@pytest.mark.parametrize("kind", ["noise", "text"])
def test_sauvola_page_bit_exact(cuda, shape, window, kind):
	img = synth.make_page(shape[1], shape[0], seed=window, kind=kind)
	want = exact.binarize(img, ("sauvola", window))
	got = ops.sauvola(ops.to_device_image(img, cuda), window).cpu().numpy()
	assert np.array_equal(got, want)


# the warp kernel of windows <= 15 (csrc/og_sauvola_warp.cu): every window, 8 pixels per lane (8-byte aligned
# rows of >= 256 columns), 4 pixels per lane (4-byte aligned), unaligned rows, several stripes and several
# runs of rows, widths and heights below the window, views into a larger image
@pytest.mark.parametrize("shape,window", [((130, 256), 3), ((77, 1000), 5), ((200, 264), 9), ((320, 488), 11), ((61, 744), 13),
	((260, 960), 15), ((50, 252), 13), ((90, 116), 5), ((33, 253), 9), ((120, 301), 11), ((1, 1), 3), ((2, 3), 15), ((3, 40), 15),
	((1, 300), 7), ((300, 1), 7), ((17, 6), 13), ((9, 257), 15)])
def test_sauvola_warp_kernel_paths(cuda, shape, window):
	img = synth.make_page(max(shape[1], 16), max(shape[0], 16), seed=shape[0] + window, kind="text")[:shape[0], :shape[1]]
	img = np.ascontiguousarray(img)
	want = exact.binarize(img, ("sauvola", window))
	got = ops.sauvola(ops.to_device_image(img, cuda), window).cpu().numpy()
	assert np.array_equal(got, want)


@pytest.mark.parametrize("x0,y0,window", [(0, 0, 15), (5, 3, 15), (8, 1, 9), (4, 2, 5), (3, 0, 3)])
def test_sauvola_on_views(cuda, x0, y0, window):
	big = synth.make_page(1100, 300, seed=7, kind="noise")
	h, w = 211, 1000
	want = exact.binarize(np.ascontiguousarray(big[y0:y0 + h, x0:x0 + w]), ("sauvola", window))
	dbig = ops.to_device_image(big, cuda)
	dst_big = torch.zeros_like(dbig)
	got = ops.sauvola(dbig[y0:y0 + h, x0:x0 + w], window, dst=dst_big[y0:y0 + h, x0:x0 + w])
	assert np.array_equal(got.cpu().numpy(), want)
	dst_big[y0:y0 + h, x0:x0 + w] = 0
	assert int(dst_big.count_nonzero().item()) == 0  # nothing written outside the view


def test_sauvola_rejects_even_window(cuda):
	img = ops.to_device_image(np.zeros((32, 32), np.uint8), cuda)
	with pytest.raises(ValueError):
		ops.sauvola(img, 16)
	with pytest.raises(ValueError):
		ops.sauvola(img, 1)


@pytest.mark.parametrize("kind", ["noise", "text", "flat", "two"])
def test_otsu_page_bit_exact(cuda, kind):
	if kind == "flat":
		img = np.full((100, 130), 77, np.uint8)
	elif kind == "two":
		img = np.zeros((100, 130), np.uint8)
		img[:, 50:] = 200
	else:
		img = synth.make_page(500, 300, seed=2, kind=kind)
	want = exact.binarize(img, ("otsu",))
	dst, t = ops.otsu(ops.to_device_image(img, cuda))
	assert np.array_equal(dst.cpu().numpy(), want)
	try:
		tw = exact.threshold_otsu(img)
	except ValueError:
		tw = 128
	assert int(t.item()) == tw


@pytest.mark.parametrize("mode", ["otsu", "sauvola"])
def test_strip_binarizers_bit_exact(cuda, mode):
	width, height, cell = 600, 800, 25
	page = synth.make_page(width, height, seed=1, kind="text", dpi=300)
	grid = synth.make_grid(width, height, cell, seed=1)
	dense = exact.expand_map(grid, cell)
	lines = synth.make_lines(dense.shape[1], dense.shape[0], 60, seed=1, dpi=300)
	g = torch.from_numpy(grid).to(cuda)
	src = ops.to_device_image(page, cuda)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines])
	plan = ops.line_steps(g, cell, ops.lines_to_device(la, cuda), len(lines))
	offs = plan.pix_offsets.cpu().numpy()
	gray = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=cuda)
	ops.extract_lines(src, g, cell, plan, gray)
	m = ops.OG_BIN_OTSU if mode == "otsu" else ops.OG_BIN_SAUVOLA
	out = ops.binarize_strips(gray, plan, m, window=15).cpu().numpy()
	spec = ("otsu",) if mode == "otsu" else ("sauvola", 15)
	for i, ln in enumerate(lines):
		strip = exact.dewarped_line_image(page, dense, ln["p"], ln["right"], ln["up"], 48)
		want = exact.binarize(strip, spec)
		assert np.array_equal(out[offs[i]:offs[i + 1]].reshape(want.shape), want), i


@pytest.mark.parametrize("M,C,shape", [(1, 4, (240, 128)), (5, 4, (241, 127)), (3, 6, (50, 70))])
def test_argmax_and_masks_bit_exact(cuda, M, C, shape):
	probs = synth.make_probabilities(M, shape[1], shape[0], C, seed=M)
	want = exact.argmax_labels(probs)
	values = list(range(1, C))
	labels, masks = ops.argmax_labels(torch.from_numpy(probs).to(cuda), values)
	assert np.array_equal(labels.cpu().numpy(), want)
	assert np.array_equal(masks.cpu().numpy(), exact.class_masks(want, values))
	m2 = ops.class_masks(ops.to_device_image(want, cuda), values)
	assert np.array_equal(m2.cpu().numpy(), exact.class_masks(want, values))


def test_argmax_ties_first_wins(cuda):
	probs = np.zeros((2, 8, 8, 4), np.float32)
	probs[..., 2] = 0.5
	probs[..., 3] = 0.5
	labels, _ = ops.argmax_labels(torch.from_numpy(probs).to(cuda))
	assert (labels.cpu().numpy() == 2).all()


def test_warp_affine_bit_exact(cuda):
	page = synth.make_page(600, 800, seed=4, kind="noise")
	src_tri = np.array([[100.3, 400.2], [520.9, 404.7], [99.1, 371.5]])
	width = int(np.ceil(np.linalg.norm(src_tri[1] - src_tri[0])))
	dst_tri = np.array([[0, 47], [width, 47], [0, 0]])
	M = exact.affine_from_points(src_tri, dst_tri)
	want = exact.warp_affine_linear(page, M, width, 48, border=200)
	got = ops.warp_affine(ops.to_device_image(page, cuda), M, width, 48, border_value=200).cpu().numpy()
	assert np.array_equal(got, want)


def test_tiled_label_assembly_is_write_inner(cuda):
	"""predict.py:275-285: per tile argmax + Tile.write_inner, against the NumPy expressions of the reference."""
	from origami_b200.core.predict import Tiles
	full = (1280, 2400)  # (w, h) as the reference's full_size
	rng = np.random.default_rng(3)
	want = np.empty((full[1], full[0]), np.uint8)
	labels = ops.alloc_image(full[1], full[0], cuda)
	labels.fill_(255)
	tiles = list(Tiles((512, 768), beta0=50)(full))
	assert len(tiles) > 6
	for k, tile in enumerate(tiles):
		x0, y0, x1, y1 = tile.outer
		pr = [rng.random((1, y1 - y0, x1 - x0, 4), dtype=np.float32) for _ in range(1 + k % 3)]
		tile.write_inner(want, np.argmax(np.sum(pr, axis=0).squeeze(), axis=-1).astype(np.uint8))
		tile.write_inner_argmax(labels, pr)
	assert np.array_equal(labels.cpu().numpy(), want)
	pr6 = rng.random((2, 40, 50, 6), dtype=np.float32)  # another class count: the general path
	small = ops.alloc_image(40, 50, cuda)
	ops.argmax_tile(torch.from_numpy(pr6).to(cuda), (0, 0, 50, 40), (0, 0), small)
	assert np.array_equal(small.cpu().numpy(), exact.argmax_labels(pr6))
