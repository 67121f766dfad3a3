"""
Thin torch-tensor front end of the C ABI: allocates outputs with torch, passes
`tensor.data_ptr()` and the current CUDA stream to liborigami_b200.so.  torch is used for
device memory and streams only; every computation happens in the hand-written kernels.
"""

import os

import numpy as np
import torch

from . import _lib
from ._lib import LINE_DTYPE, OG_BIN_NONE, OG_BIN_OTSU, OG_BIN_SAUVOLA, OG_INTER_LINEAR, OG_INTER_NEAREST

OG_FLAG_STAGE_SMEM = 0x100
OG_FLAG_MAP_GATHER = 0x100  # og_extract_lines: sample through the dense map with the map-gather engines

PITCH_ALIGN = 128


def _require_cuda():
	if not torch.cuda.is_available():
		raise _lib.OrigamiB200Error("origami_b200 needs a CUDA device (there is no CPU fallback)")


def _stream():
	return torch.cuda.current_stream().cuda_stream


def _call(name, anchor, spec):
	"""Runs a library call on the device that owns `anchor` (a CUDA tensor): that device is made
	current for the duration of the call and the work goes to ITS current stream -- the kernels
	launch on whatever device is current, so a tensor on cuda:1 with cuda:0 current would otherwise
	be handed to the wrong device.  `spec` is (list of other tensors that must live on the same
	device, builder) where builder(stream) returns the argument tuple."""
	others, build = spec
	dev = anchor.device
	for t in others:
		if t is not None and t.device != dev:
			raise ValueError("origami_b200: arguments live on different devices (%s and %s)" % (dev, t.device))
	if dev.index is not None and dev.index != torch.cuda.current_device():
		with torch.cuda.device(dev):
			return _lib.call(name, *build(torch.cuda.current_stream(dev).cuda_stream))
	return _lib.call(name, *build(torch.cuda.current_stream(dev).cuda_stream))


def _ptr(t):
	return 0 if t is None else t.data_ptr()


def _check_u8_image(t, name):
	if t.dtype != torch.uint8 or t.dim() != 2 or not t.is_cuda or t.stride(1) != 1:
		raise ValueError("%s must be a 2-D uint8 CUDA tensor with unit column stride" % name)


def alloc_image(h, w, device=None):
	"""(h, w) uint8 view into a buffer whose row pitch is a multiple of 128 bytes, so the
	kernels can use aligned vector stores for any width."""
	_require_cuda()
	pitch = (w + PITCH_ALIGN - 1) // PITCH_ALIGN * PITCH_ALIGN
	buf = torch.empty((h, pitch), dtype=torch.uint8, device=device or "cuda")
	return buf[:, :w]


_PINNED = {}


def pinned_staging(nbytes):
	"""A reusable page-locked host buffer of at least nbytes (host<->device copies go through
	pinned memory so they run at PCIe speed and can be asynchronous)."""
	buf = _PINNED.get("u8")
	if buf is None or buf.numel() < nbytes:
		buf = torch.empty((max(nbytes, 1 << 20),), dtype=torch.uint8).pin_memory()
		_PINNED["u8"] = buf
	return buf[:nbytes]


_COPY_POOL = None
# four threads and four row bands are the measured optimum on the 16-core hosts of the B200 boxes
# (profiles/r2_upload_paths.txt: 35 MB page in 1.5 ms; 8 threads 3.4 ms, 1 thread 3.5 ms)
_COPY_THREADS = max(2, min(4, (os.cpu_count() or 4) // 2))


def parallel_copy(dst, src):
	"""dst[:n] = src for large contiguous uint8 arrays, split over a few threads (NumPy releases the
	GIL while copying; one thread reaches only a fraction of the host memory bandwidth)."""
	global _COPY_POOL
	n = src.shape[0]
	if n < (2 << 20):
		dst[:n] = src
		return
	if _COPY_POOL is None:
		import concurrent.futures
		_COPY_POOL = concurrent.futures.ThreadPoolExecutor(max_workers=_COPY_THREADS)
	step = (n + _COPY_THREADS - 1) // _COPY_THREADS
	futs = [_COPY_POOL.submit(np.copyto, dst[i:min(n, i + step)], src[i:min(n, i + step)]) for i in range(0, n, step)]
	for f in futs:
		f.result()


_STAGE_CHUNKS = 4


def to_device_image(a, device=None):
	"""Host uint8 (h, w) ndarray -> pitched device image through pinned staging memory.  The page is
	cut into a few row bands: while band k travels to the device, band k+1 is copied into the pinned
	buffer by the host threads, so the upload costs about the slower of the two instead of their sum."""
	a = np.ascontiguousarray(a, dtype=np.uint8)
	h, w = a.shape
	out = alloc_image(h, w, device)
	stage = pinned_staging(h * w)
	stage_np = stage.numpy()
	flat = a.reshape(-1)
	stream = torch.cuda.current_stream(out.device)
	rows = max(1, (h + _STAGE_CHUNKS - 1) // _STAGE_CHUNKS) if h * w >= (8 << 20) else h
	with torch.cuda.device(out.device):
		for r0 in range(0, h, rows):
			r1 = min(h, r0 + rows)
			parallel_copy(stage_np[r0 * w:r1 * w], flat[r0 * w:r1 * w])  # PIL arrays are read-only views: copy, never pin in place
			out[r0:r1].copy_(stage[r0 * w:r1 * w].view(r1 - r0, w), non_blocking=True)
	stream.synchronize()  # the staging buffer is reused by the next call
	return out


class HostCopy:
	"""A device->host copy in flight on a side stream: result() waits for it and returns the ndarray."""

	def __init__(self, t):
		dev = t.device
		self._host = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
		side = _side_stream(dev)
		side.wait_stream(torch.cuda.current_stream(dev))
		with torch.cuda.stream(side):
			self._host.copy_(t, non_blocking=True)
			self._done = torch.cuda.Event()
			self._done.record(side)
		t.record_stream(side)
		self._array = None

	def result(self):
		if self._array is None:
			self._done.synchronize()
			self._array = self._host.numpy()
		return self._array


_SIDE = {}


def _side_stream(dev):
	key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
	st = _SIDE.get(key)
	if st is None:
		st = _SIDE[key] = torch.cuda.Stream(device=dev)
	return st


def to_host(t):
	"""Device tensor -> NumPy array through page-locked memory (torch caches pinned blocks, so
	this is one DMA at PCIe speed instead of a staged pageable copy)."""
	host = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
	host.copy_(t, non_blocking=True)
	torch.cuda.current_stream(t.device).synchronize()
	return host.numpy()


def _check_grid(grid):
	if grid.dtype != torch.float32 or grid.dim() != 3 or grid.shape[2] != 2 or not grid.is_cuda or not grid.is_contiguous():
		raise ValueError("grid must be a contiguous float32 CUDA tensor of shape (gh, gw, 2)")
	return int(grid.shape[0]), int(grid.shape[1])


def expand_map(grid, cell, out=None):
	"""Grid.points("full") (reference origami/core/dewarp.py:514-526)."""
	_require_cuda()
	gh, gw = _check_grid(grid)
	if out is None:
		out = torch.empty((gh * cell, gw * cell, 2), dtype=torch.float32, device=grid.device)
	_call("og_expand_map", grid, ([out], lambda st: (grid.data_ptr(), gh, gw, cell, out.data_ptr(), st)))
	return out


def remap_page(src, grid, cell, dst=None, map_out=None, scale=(1.0, 1.0), interpolation=OG_INTER_LINEAR,
		want_image=True):
	"""Dewarper.dewarped / dewarp_image (reference origami/core/dewarp.py:608-627) fused with
	the map expansion.  Returns (dst, map_out)."""
	_require_cuda()
	gh, gw = _check_grid(grid)
	H = W = 0
	if want_image:
		_check_u8_image(src, "src")
		H, W = int(src.shape[0]), int(src.shape[1])
		if dst is None:
			dst = alloc_image(gh * cell, gw * cell, grid.device)
		_check_u8_image(dst, "dst")
	_call("og_remap_page", grid, ([src if want_image else None, dst if want_image else None, map_out], lambda st: (
		_ptr(src) if want_image else 0, H, W, src.stride(0) if want_image else 0,
		grid.data_ptr(), gh, gw, cell,
		_ptr(dst) if want_image else 0, dst.stride(0) if want_image else 0, _ptr(map_out),
		float(np.float32(scale[0])), float(np.float32(scale[1])), int(interpolation), st)))
	return dst, map_out


def inverse_yx(grid, cell, pts_yx):
	"""Grid.inverse_yx (reference origami/core/dewarp.py:561-567) on an (n, 2) float64 tensor."""
	_require_cuda()
	gh, gw = _check_grid(grid)
	pts = pts_yx.contiguous()
	if pts.dtype != torch.float64 or pts.dim() != 2 or pts.shape[1] != 2:
		raise ValueError("pts_yx must be float64 (n, 2)")
	out = torch.empty_like(pts)
	_call("og_inverse_yx", grid, ([pts], lambda st: (grid.data_ptr(), gh, gw, cell, pts.data_ptr(), pts.shape[0], out.data_ptr(), st)))
	return out


def pack_lines(records):
	"""records: iterable of dicts with p, right, up and optional cp, cright, xres, yres, xsteps,
	ysteps -> structured ndarray of og_line."""
	records = list(records)
	n = len(records)
	a = np.zeros(n, dtype=LINE_DTYPE)
	if n == 0:
		return a
	a["p"] = np.array([r["p"] for r in records], dtype=np.float64).reshape(n, 2)
	a["right"] = np.array([r["right"] for r in records], dtype=np.float64).reshape(n, 2)
	a["up"] = np.array([r["up"] for r in records], dtype=np.float64).reshape(n, 2)
	a["cp"] = np.array([r.get("cp", r["p"]) for r in records], dtype=np.float64).reshape(n, 2)
	a["cright"] = np.array([r.get("cright", r["right"]) for r in records], dtype=np.float64).reshape(n, 2)
	a["xres"] = [r.get("xres", 1.0) for r in records]
	a["yres"] = [r.get("yres", 1.0) for r in records]
	a["xsteps"] = [r.get("xsteps") or 0 for r in records]
	a["ysteps"] = [r.get("ysteps") or 0 for r in records]
	return a


def pack_line_arrays(p, right, up, xsteps=0, ysteps=0, xres=1.0, yres=1.0):
	"""Vectorised pack_lines for un-clipped lines: p, right, up are (n, 2) float64 arrays."""
	def table(v):  # a list of n (2,) arrays is converted once (np.concatenate is the cheap way through such a list)
		if isinstance(v, np.ndarray):
			return np.asarray(v, dtype=np.float64).reshape(-1, 2)
		return np.concatenate(v).astype(np.float64, copy=False).reshape(-1, 2) if len(v) else np.zeros((0, 2))
	p, right, up = table(p), table(right), table(up)
	a = np.zeros(p.shape[0], dtype=LINE_DTYPE)
	a["p"] = p
	a["cp"] = p
	a["right"] = right
	a["cright"] = right
	a["up"] = up
	a["xres"] = xres
	a["yres"] = yres
	a["xsteps"] = xsteps
	a["ysteps"] = ysteps
	return a


def lines_to_device(lines_np, device=None):
	_require_cuda()
	raw = lines_np.view(np.uint8).reshape(-1)
	stage = torch.empty((raw.shape[0],), dtype=torch.uint8, pin_memory=True)
	stage.numpy()[...] = raw
	return stage.to(device or "cuda", non_blocking=True)


class LinePlan:
	"""Device-side step counts and offsets of a batch of lines (output of og_line_steps)."""

	def __init__(self, lines_dev, n, steps, pix_offsets, chunk_offsets):
		self.lines = lines_dev
		self.n = n
		self.steps = steps
		self.pix_offsets = pix_offsets
		self.chunk_offsets = chunk_offsets


def line_steps(grid, cell, lines_dev, n):
	"""Step counts of Line.warped_grid (reference origami/core/block.py:226-236) + offsets.  The three
	device arrays of the plan live in one buffer ([pix offsets | steps | chunk offsets]), so that the
	host can fetch offsets and steps with a single copy (plan_to_host)."""
	_require_cuda()
	gh, gw = _check_grid(grid)
	dev = grid.device
	o_steps = 8 * (n + 1)
	o_chunks = o_steps + 8 * max(n, 1)
	buf = torch.empty((o_chunks + 4 * (n + 2),), dtype=torch.uint8, device=dev)
	pix = buf[:o_steps].view(torch.int64)
	steps = buf[o_steps:o_chunks].view(torch.int32).view(max(n, 1), 2)
	chunks = buf[o_chunks:].view(torch.int32)  # n + 1 prefix sums + the work ticket
	_call("og_line_steps", grid, ([lines_dev], lambda st: (grid.data_ptr(), gh, gw, cell, lines_dev.data_ptr(), n,
		steps.data_ptr(), pix.data_ptr(), chunks.data_ptr(), st)))
	plan = LinePlan(lines_dev, n, steps, pix, chunks)
	plan.buf = buf
	return plan


def plan_to_host(plan):
	"""(steps (n, 2) int32, pix_offsets (n + 1) int64) of a plan made by line_steps: one device->host copy."""
	n = plan.n
	nbytes = 8 * (n + 1) + 8 * max(n, 1)
	host = torch.empty((nbytes,), dtype=torch.uint8, pin_memory=True)
	host.copy_(plan.buf[:nbytes], non_blocking=True)
	torch.cuda.current_stream(plan.buf.device).synchronize()
	a = host.numpy()
	return a[8 * (n + 1):].view(np.int32).reshape(max(n, 1), 2)[:n], a[:8 * (n + 1)].view(np.int64)


def _check_map(dense_map, gh, gw, cell):
	if dense_map is None:
		return 0
	if (dense_map.dtype != torch.float32 or not dense_map.is_cuda or not dense_map.is_contiguous()
			or tuple(dense_map.shape) != (gh * cell, gw * cell, 2)):
		raise ValueError("dense_map must be the contiguous float32 CUDA tensor (gh*cell, gw*cell, 2) of this grid")
	return dense_map.data_ptr()


def warped_grids(grid, cell, plan, out, dense_map=None):
	"""Line.warped_grid for every line; out: float32 buffer of 2*capacity elements."""
	gh, gw = _check_grid(grid)
	_call("og_warped_grids", grid, ([dense_map, plan.lines, plan.steps, out], lambda st: (
		grid.data_ptr(), gh, gw, cell, _check_map(dense_map, gh, gw, cell), plan.lines.data_ptr(), plan.n,
		plan.steps.data_ptr(), plan.pix_offsets.data_ptr(), plan.chunk_offsets.data_ptr(),
		out.data_ptr(), out.numel() // 2, st)))
	return out


def _binarize_args(binarize):
	"""None | ("otsu",) | ("sauvola", window) | OG_BIN_* -> (mode, window)."""
	if binarize is None or binarize == OG_BIN_NONE:
		return OG_BIN_NONE, 0
	if isinstance(binarize, int):
		return int(binarize), 15
	if binarize[0] == "otsu":
		return OG_BIN_OTSU, 0
	if binarize[0] == "sauvola":
		return OG_BIN_SAUVOLA, int(binarize[1]) if len(binarize) > 1 else 15
	raise ValueError("unknown binarizer %r" % (binarize,))


def extract_lines(src, grid, cell, plan, out, bbox=None, dense_map=None, binarize=None, gray_out=None, thresholds=None,
		map_gather=False):
	"""Line.image(dewarped=True, binarizer=...) for every line into the packed uint8 buffer `out`
	(reference origami/core/block.py:129-143): gray strips, or binarized ones.  binarize: None |
	("otsu",) | ("sauvola", window); Sauvola needs its gray input whole, so a gray scratch buffer is
	allocated when gray_out is not given.  With dense_map the nodes are gathered from it (fastest),
	without it they are recomputed on chip (no dense map needed at all)."""
	gh, gw = _check_grid(grid)
	_check_u8_image(src, "src")
	mode, window = _binarize_args(binarize)
	if mode == OG_BIN_SAUVOLA and gray_out is None:
		gray_out = torch.empty_like(out)
	if map_gather:
		mode |= OG_FLAG_MAP_GATHER
	cap = out.numel() if gray_out is None else min(out.numel(), gray_out.numel())
	_call("og_extract_lines", grid, ([src, dense_map, plan.lines, plan.steps, out, bbox, gray_out, thresholds], lambda st: (
		src.data_ptr(), int(src.shape[0]), int(src.shape[1]), src.stride(0),
		grid.data_ptr(), gh, gw, cell, _check_map(dense_map, gh, gw, cell), plan.lines.data_ptr(), plan.n,
		plan.steps.data_ptr(), plan.pix_offsets.data_ptr(), plan.chunk_offsets.data_ptr(),
		out.data_ptr(), cap, _ptr(bbox), mode, window, _ptr(gray_out), _ptr(thresholds), st)))
	return out


def binarize_strips(gray, plan, mode, window=15, out=None, thresholds=None):
	if out is None:
		out = torch.empty_like(gray)
	_call("og_binarize_strips", gray, ([plan.steps, out, thresholds], lambda st: (
		gray.data_ptr(), plan.steps.data_ptr(), plan.pix_offsets.data_ptr(),
		plan.chunk_offsets.data_ptr(), plan.n, int(mode), int(window), out.data_ptr(), min(gray.numel(), out.numel()),
		_ptr(thresholds), st)))
	return out


def pack_strips(strips, plan, height, w_max, min_w=6, min_h=6, pad_value=0, batch=None):
	"""The recogniser's input batch, assembled on the device (reference origami/batch/detect/ocr.py:
	238-268): strips of at least min_w x min_h pixels, in line order, right-padded to w_max.
	Returns (batch (n, height, w_max) uint8 -- the first n_kept slots are filled --, widths (n) int32,
	kept_index (n) int32, n_kept (1) int32), all on the device."""
	_require_cuda()
	n = plan.n
	dev = strips.device
	if batch is None:
		batch = torch.empty((max(n, 1), height, w_max), dtype=torch.uint8, device=dev)
	meta = torch.empty((3 * max(n, 1) + 1,), dtype=torch.int32, device=dev)
	slots, widths, kept, n_kept = meta[:n], meta[max(n, 1):max(n, 1) + n], meta[2 * max(n, 1):2 * max(n, 1) + n], meta[3 * max(n, 1):]
	_call("og_pack_strips", strips, ([plan.steps, batch], lambda st: (
		strips.data_ptr(), plan.steps.data_ptr(), plan.pix_offsets.data_ptr(), n, strips.numel(),
		int(min_w), int(min_h), int(pad_value), int(height), int(w_max),
		batch.data_ptr(), slots.data_ptr(), widths.data_ptr(), kept.data_ptr(), n_kept.data_ptr(), st)))
	return batch, widths, kept, n_kept


def pack_bits(src, n=None, out=None):
	"""Binarized bytes -> one bit per pixel on the device (numpy.packbits order): src a flat uint8 CUDA
	tensor, n pixels (all of it by default); returns a uint8 tensor of (n + 7) // 8 bytes."""
	_require_cuda()
	if src.dtype != torch.uint8 or not src.is_cuda or not src.is_contiguous():
		raise ValueError("src must be a contiguous uint8 CUDA tensor")
	n = src.numel() if n is None else int(n)
	if n > src.numel():
		raise ValueError("n exceeds the tensor")
	if out is None:
		out = torch.empty(((n + 7) // 8,), dtype=torch.uint8, device=src.device)
	_call("og_pack_bits", src, ([out], lambda st: (src.data_ptr(), n, out.data_ptr(), st)))
	return out


def sauvola(src, window=15, k=0.2, r=127.5, dst=None):
	"""(src > threshold_sauvola(src, window)) * 255 on a device image."""
	_require_cuda()
	_check_u8_image(src, "src")
	h, w = int(src.shape[0]), int(src.shape[1])
	if dst is None:
		dst = alloc_image(h, w, src.device)
	_call("og_sauvola", src, ([dst], lambda st: (src.data_ptr(), h, w, src.stride(0), int(window), float(k), float(r),
		dst.data_ptr(), dst.stride(0), st)))
	return dst


def otsu(src, dst=None, hist_scratch=None, threshold=None):
	"""(src > threshold_otsu(src)) * 255 on a device image; returns (dst, threshold tensor)."""
	_require_cuda()
	_check_u8_image(src, "src")
	h, w = int(src.shape[0]), int(src.shape[1])
	if dst is None:
		dst = alloc_image(h, w, src.device)
	if hist_scratch is None:
		hist_scratch = torch.empty((256,), dtype=torch.int32, device=src.device)
	if threshold is None:
		threshold = torch.empty((1,), dtype=torch.int32, device=src.device)
	_call("og_otsu", src, ([dst, hist_scratch, threshold], lambda st: (src.data_ptr(), h, w, src.stride(0), dst.data_ptr(), dst.stride(0),
		hist_scratch.data_ptr(), threshold.data_ptr(), st)))
	return dst, threshold


def sample_labels(labels, page_shape, grid, cell, plan, n_classes, dense_map=None):
	"""ConfidenceSampler counts (reference origami/batch/detect/lines.py:34-54)."""
	gh, gw = _check_grid(grid)
	_check_u8_image(labels, "labels")
	counts = torch.empty((max(plan.n, 1), n_classes), dtype=torch.int64, device=grid.device)
	overflow = torch.empty((max(plan.n, 1),), dtype=torch.int64, device=grid.device)
	_call("og_sample_labels", grid, ([labels, dense_map, plan.lines, plan.steps], lambda st: (
		labels.data_ptr(), int(labels.shape[0]), int(labels.shape[1]), labels.stride(0),
		int(page_shape[0]), int(page_shape[1]), grid.data_ptr(), gh, gw, cell, _check_map(dense_map, gh, gw, cell),
		plan.lines.data_ptr(), plan.n,
		plan.steps.data_ptr(), plan.chunk_offsets.data_ptr(), n_classes, counts.data_ptr(), overflow.data_ptr(), st)))
	return counts[:plan.n], overflow[:plan.n]


def argmax_labels(probs, mask_values=None):
	"""probs (M, h, w, C) float32 -> (labels (h, w) uint8, masks (n_mask, h, w) uint8 | None)."""
	_require_cuda()
	if probs.dtype != torch.float32 or probs.dim() != 4 or not probs.is_cuda or not probs.is_contiguous():
		raise ValueError("probs must be a contiguous float32 CUDA tensor (M, h, w, C)")
	M, h, w, C = (int(v) for v in probs.shape)
	labels = torch.empty((h, w), dtype=torch.uint8, device=probs.device)
	masks = None
	mv = None
	n_mask = 0
	if mask_values is not None and len(mask_values) > 0:
		mv = np.ascontiguousarray(mask_values, dtype=np.uint8)
		n_mask = int(mv.shape[0])
		masks = torch.empty((n_mask, h, w), dtype=torch.uint8, device=probs.device)
	_call("og_argmax_labels", probs, ([], lambda st: (probs.data_ptr(), M, C, h, w, labels.data_ptr(), _ptr(masks),
		mv.ctypes.data if mv is not None else 0, n_mask, st)))
	return labels, masks


def argmax_tile(probs, inner, outer_origin, labels):
	"""Tile.write_inner (reference origami/core/predict.py:60-63) fused with the argmax: probs (M, th,
	tw, C) of one tile; inner = (x0, y0, x1, y1) in page coordinates, outer_origin = (x, y) of the
	tile's top-left corner; the inner rectangle of the tile's labels lands in `labels` (device image)."""
	_require_cuda()
	if probs.dtype != torch.float32 or probs.dim() != 4 or not probs.is_cuda or not probs.is_contiguous():
		raise ValueError("probs must be a contiguous float32 CUDA tensor (M, th, tw, C)")
	_check_u8_image(labels, "labels")
	M, th, tw, C = (int(v) for v in probs.shape)
	x0, y0, x1, y1 = (int(v) for v in inner)
	dx, dy = x0 - int(outer_origin[0]), y0 - int(outer_origin[1])
	_call("og_argmax_tile", probs, ([labels], lambda st: (probs.data_ptr(), M, C, th, tw, dx, dy, x1 - x0, y1 - y0,
		labels.data_ptr(), int(labels.shape[0]), int(labels.shape[1]), labels.stride(0), x0, y0, st)))
	return labels


def class_masks(labels, mask_values):
	_require_cuda()
	_check_u8_image(labels, "labels")
	mv = np.ascontiguousarray(mask_values, dtype=np.uint8)
	h, w = int(labels.shape[0]), int(labels.shape[1])
	masks = torch.empty((int(mv.shape[0]), h, w), dtype=torch.uint8, device=labels.device)
	_call("og_class_masks", labels, ([], lambda st: (labels.data_ptr(), h, w, labels.stride(0), masks.data_ptr(), mv.ctypes.data,
		int(mv.shape[0]), st)))
	return masks


def warp_affine(src, M, width, height, border_value=0):
	"""cv2.warpAffine(src, M, (width, height), INTER_LINEAR, BORDER_CONSTANT, border_value)."""
	_require_cuda()
	_check_u8_image(src, "src")
	Mh = np.ascontiguousarray(M, dtype=np.float64).reshape(6)
	dst = alloc_image(height, width, src.device)
	_call("og_warp_affine", src, ([], lambda st: (src.data_ptr(), int(src.shape[0]), int(src.shape[1]), src.stride(0),
		Mh.ctypes.data, dst.data_ptr(), width, height, dst.stride(0), int(border_value), st)))
	return dst


def column_psd_quantile(pixels, quantile=0.9):
	"""np.quantile(scipy.signal.periodogram(pixels, axis=0)[1], quantile, axis=0) for a 2-D uint8 or
	float32 CUDA tensor with unit column stride (reference origami/batch/detect/layout.py:926-930)."""
	_require_cuda()
	if pixels.dim() != 2 or not pixels.is_cuda or pixels.stride(1) != 1 or pixels.dtype not in (torch.uint8, torch.float32):
		raise ValueError("pixels must be a 2-D uint8 or float32 CUDA tensor with unit column stride")
	h, w = int(pixels.shape[0]), int(pixels.shape[1])
	out = torch.empty((w,), dtype=torch.float32, device=pixels.device)
	_call("og_column_psd_quantile", pixels, ([], lambda st: (pixels.data_ptr(), 1 if pixels.dtype == torch.uint8 else 0, h, w,
		pixels.stride(0) * pixels.element_size(), float(quantile), out.data_ptr(), st)))
	return out


def region_psd_quantile(pixels, boxes, transposed=False, quantile=0.9):
	"""column_psd_quantile for many rectangles of one device image in one launch: boxes (m, 4) int
	(x0, y0, x1, y1), already clipped to the image; transposed: a flag or one flag per box -- the series
	run along the rows of the box (crop.transpose() in the reference).  Returns a list of m float32
	NumPy arrays (one value per column of the box, or per row when transposed)."""
	_require_cuda()
	if pixels.dim() != 2 or not pixels.is_cuda or pixels.stride(1) != 1 or pixels.dtype not in (torch.uint8, torch.float32):
		raise ValueError("pixels must be a 2-D uint8 or float32 CUDA tensor with unit column stride")
	H, W = int(pixels.shape[0]), int(pixels.shape[1])
	boxes = np.asarray(boxes, dtype=np.int64).reshape(-1, 4)
	m = boxes.shape[0]
	if m == 0:
		return []
	tr = np.broadcast_to(np.asarray(transposed, dtype=bool), (m,))
	w, h = boxes[:, 2] - boxes[:, 0], boxes[:, 3] - boxes[:, 1]
	if (boxes[:, 0] < 0).any() or (boxes[:, 1] < 0).any() or (boxes[:, 2] > W).any() or (boxes[:, 3] > H).any() or (w < 0).any() or (h < 0).any():
		raise ValueError("boxes must lie inside the image")
	count = np.where((w > 0) & (h > 0), np.where(tr, h, w), 0)
	length = np.where(tr, w, h)
	offs = np.concatenate([[0], np.cumsum(count)]).astype(np.int64)
	total = int(offs[-1])
	if total == 0:
		return [np.zeros(0, np.float32) for _ in range(m)]
	if total >= 2 ** 31:
		raise ValueError("too many series")
	max_len = int(length[count > 0].max())
	rec = np.stack([boxes[:, 0], boxes[:, 1], w, h, tr.astype(np.int64)], axis=1).astype(np.int32)
	meta = torch.from_numpy(np.concatenate([rec.reshape(-1), offs.astype(np.int32)])).to(pixels.device)
	out = torch.empty((total,), dtype=torch.float32, device=pixels.device)
	_call("og_region_psd_quantile", pixels, ([meta], lambda st: (pixels.data_ptr(), 1 if pixels.dtype == torch.uint8 else 0, H, W,
		pixels.stride(0) * pixels.element_size(), meta.data_ptr(), meta.data_ptr() + 4 * 5 * m, m, total, max_len, float(quantile),
		out.data_ptr(), st)))
	flat = out.cpu().numpy()
	return [flat[offs[i]:offs[i + 1]] for i in range(m)]


def polygon_mask(verts_xy, w, h, device=None, image=None, background=255):
	"""Mask.__init__ / Mask._extract (reference origami/core/mask.py:8-74): verts_xy (k, 2) ring relative
	to the mask origin.  Returns the (h, w) uint8 mask (255 inside); with `image` (a device image of
	that size) the pixels outside the polygon are set to background in place as well."""
	_require_cuda()
	v = np.ascontiguousarray(verts_xy, dtype=np.float64).reshape(-1, 2)
	if len(v) > 1 and np.array_equal(v[0], v[-1]):
		v = v[:-1]  # closed ring: the closing edge is implied
	dev = image.device if image is not None else torch.device(device or "cuda")
	vd = torch.from_numpy(v).to(dev)
	mask = alloc_image(h, w, dev)
	if image is not None:
		_check_u8_image(image, "image")
	_call("og_polygon_mask", mask, ([vd, image], lambda st: (vd.data_ptr(), int(v.shape[0]), int(w), int(h),
		mask.data_ptr(), mask.stride(0), _ptr(image), image.stride(0) if image is not None else 0, int(background), st)))
	return mask


def label_mask_ne(labels, background):
	"""(labels != background) * 255 on a device label map."""
	_require_cuda()
	_check_u8_image(labels, "labels")
	h, w = int(labels.shape[0]), int(labels.shape[1])
	out = alloc_image(h, w, labels.device)
	_call("og_label_mask_ne", labels, ([], lambda st: (labels.data_ptr(), h, w, labels.stride(0), int(background), out.data_ptr(), out.stride(0), st)))
	return out


def or_dilate3(binary, mask=None):
	"""float32 (h, w): binary != 0 OR 3x3-dilated mask != 0."""
	_require_cuda()
	_check_u8_image(binary, "binary")
	h, w = int(binary.shape[0]), int(binary.shape[1])
	if mask is not None:
		_check_u8_image(mask, "mask")
		if tuple(mask.shape) != (h, w):
			raise ValueError("mask shape differs from the binary image")
	out = torch.empty((h, w), dtype=torch.float32, device=binary.device)
	_call("og_or_dilate3", binary, ([mask], lambda st: (binary.data_ptr(), binary.stride(0), _ptr(mask), mask.stride(0) if mask is not None else 0,
		h, w, out.data_ptr(), st)))
	return out
