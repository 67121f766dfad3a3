"""
Multi-page pipeline for the page-geometry path: while the GPU works on page i, the host
packs and uploads page i+1 and the results of page i-1 travel back -- pinned staging
buffers, one CUDA stream per in-flight page, no synchronisation inside a page.

The reference processes its page queue one page at a time inside `Processor.process`
(origami/batch/core/processor.py:396-441); every page is independent, so nothing but host
scheduling changes here.  Per page the work is what `Page(..., grid).dewarped` plus
`LineExtractor(...)(lines)` do: fused map expansion + remap, line step counts, strip
extraction and optional strip binarization, with bit-identical results.
"""

import collections
import collections.abc
import weakref

import numpy as np
import torch

from ... import _lib, ops
from ...core import binarize as _binarize

PageResult = collections.namedtuple("PageResult", ["dewarped", "strips", "steps"])
# strips of a page handed over on the device: batch (n_kept, line_height, w_max) uint8 CUDA tensor (valid
# until the slot is reused, like the host views), widths and line indices of the kept strips (NumPy)
DeviceBatch = collections.namedtuple("DeviceBatch", ["batch", "widths", "line_index"])

from ...core.block import PackedStripList, StripList  # noqa: E402  (re-exported: results of PagePipeline use them too)


_parallel_copy = ops.parallel_copy


class _Slot:
	"""Buffers of one in-flight page (reused across pages; grown on demand)."""

	def __init__(self, device):
		self.device = device
		self.stream = torch.cuda.Stream(device=device)
		self.done = None
		self.h_in = self.h_out = None
		self.d_in = self.d_out = self.d_gray = self.d_chunks = self.d_map = None
		self.d_strips = self.d_batch = self.d_slots = None
		self.meta = None
		self.bytes = (0, 0)

	@staticmethod
	def _grow(t, nbytes, **kw):
		if t is None or t.numel() < nbytes:
			# 25 % headroom: pages of a batch differ a little in size, and growing a pinned or
			# device buffer in the middle of a run synchronises the device
			return torch.empty((max(nbytes + nbytes // 4, 1 << 20),), dtype=torch.uint8, **kw)
		return t


_PINNED = {}  # data pointer -> pinned tensor backing an array handed out by pinned_page()


def pinned_page(pixels_or_shape):
	"""A page-locked uint8 (H, W) array (a copy of `pixels`, or uninitialised for a shape) that an
	image decoder can fill in place.  Pages that live in such arrays are uploaded straight from
	them; any other array is first copied into the pipeline's own pinned staging buffer."""
	if isinstance(pixels_or_shape, np.ndarray):
		src = np.ascontiguousarray(pixels_or_shape, dtype=np.uint8)
		shape = src.shape
	else:
		src, shape = None, tuple(pixels_or_shape)
	t = torch.empty(shape, dtype=torch.uint8, pin_memory=True)
	a = t.numpy()
	if src is not None:
		a[...] = src
	key = a.ctypes.data
	_PINNED[key] = t
	weakref.finalize(a, _PINNED.pop, key, None)  # the registry entry (and the pinned block) goes with the array
	return a


def _estimate_xsteps(grid, cell, desc):
	"""Upper estimate of the strip widths og_line_steps will derive (Line.warped_grid's rough 2x2
	grid, block.py:226-236), from a float64 bilinear evaluation of the sparse grid on the host.
	It only sizes the strip buffer and the device->host copy: two columns of slack per line, and a
	page whose strips still do not fit is run again with the exact size."""
	xs = desc["xsteps"].astype(np.int64)
	derive = xs <= 0
	if derive.any():
		gx = grid[..., 0].astype(np.float64)
		gh, gw = gx.shape
		p, right, up = desc["p"][derive], desc["right"][derive], desc["up"][derive]
		corners = np.stack([p, p + right, up + p, up + p + right], axis=1)  # (m, 4, 2) dewarped (x, y)
		# cv2.resize replicates the border (the dense map is flat in its outer half cell) and SciPy
		# extrapolates that flat rim, so sparse coordinates are simply clamped
		u = np.clip(np.nan_to_num((corners[..., 0] + 0.5) / cell - 0.5), 0, gw - 1)
		v = np.clip(np.nan_to_num((corners[..., 1] + 0.5) / cell - 0.5), 0, gh - 1)
		with np.errstate(invalid="ignore"):
			j0 = np.minimum(np.floor(u), gw - 2).astype(np.int64)
			i0 = np.minimum(np.floor(v), gh - 2).astype(np.int64)
			fu, fv = u - j0, v - i0
			wx = (gx[i0, j0] * (1 - fu) + gx[i0, j0 + 1] * fu) * (1 - fv) + (gx[i0 + 1, j0] * (1 - fu) + gx[i0 + 1, j0 + 1] * fu) * fv
			ext = np.maximum(np.abs(wx[:, 0] - wx[:, 1]), np.abs(wx[:, 2] - wx[:, 3])) * desc["xres"][derive]
			ext = np.where(np.isfinite(ext), np.minimum(ext, 1.0e6), 0.0)
		xs[derive] = np.maximum(2, np.ceil(ext * 1.002)) + 2
	return xs


class PagePipeline:
	"""pipeline = PagePipeline(line_height=48, binarize="otsu"); for res in pipeline.run(jobs): ...

	A job is (page uint8 (H, W) ndarray, grid (gh, gw, 2) float32 ndarray, cell, lines) with
	`lines` an (n, 3, 2) float64 array or list of (p, right, up) in dewarped space (the fields
	of lines.zip), or a structured array of og_line descriptors (ops.pack_lines) when table rows
	are split into column cells.  Results come back in job order as PageResult(dewarped ndarray,
	sequence of strip ndarrays, steps (n, 2) int32 (xsteps, ysteps)).  The arrays of a result are views into a
	pinned buffer that is recycled when the NEXT result is requested: copy what must outlive that.
	With dewarped=False no dewarped page is computed (PageResult.dewarped is None): strips only."""

	def __init__(self, line_height=48, binarize="", depth=2, device=None, use_map=True, dewarped=True, hand_off="host",
			min_size=(6, 6), pad_value=0):
		if not torch.cuda.is_available():
			raise _lib.OrigamiB200Error("origami_b200 needs a CUDA device (there is no CPU fallback)")
		self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
		self.line_height = int(line_height)
		spec = _binarize.from_string(binarize).spec if binarize.strip() else None
		self.binarize = spec
		# dewarped=False: the OCR stage's configuration (origami/batch/detect/ocr.py:227-249 samples the
		# strips from the WARPED page and never touches the dewarped image): only the dense map is
		# expanded on the device, no dewarped page is produced or downloaded
		self.want_dewarped = bool(dewarped)
		self.use_map = use_map or not self.want_dewarped
		# hand_off="device": the strips stay on the GPU as the recogniser's padded input batch
		# (origami/batch/detect/ocr.py:238-268); only step counts, widths and indices travel back
		# hand_off="bits": binarized strips travel to the host as one bit per pixel (og_pack_bits) and are
		# unpacked strip by strip when they are accessed (PackedStripList)
		if hand_off not in ("host", "device", "bits"):
			raise ValueError("hand_off must be 'host', 'device' or 'bits'")
		if hand_off == "bits" and spec is None:
			raise ValueError("hand_off='bits' needs a binarizer: gray strips have more than one bit per pixel")
		self.hand_off = hand_off
		self.min_size, self.pad_value = (int(min_size[0]), int(min_size[1])), int(pad_value)
		self.slots = [_Slot(self.device) for _ in range(max(2, depth))]

	# -- one page: everything is enqueued on the slot's stream, nothing waits
	def _submit(self, slot, job, cap=None, w_cap=None):
		page, grid, cell, lines = job
		page = np.ascontiguousarray(page, dtype=np.uint8)
		grid = np.ascontiguousarray(grid, dtype=np.float32)
		H, W = page.shape
		gh, gw = grid.shape[:2]
		Hd, Wd = gh * cell, gw * cell
		if isinstance(lines, np.ndarray) and lines.dtype == ops.LINE_DTYPE:
			# ready-made og_line descriptors (e.g. table rows clipped to their columns by
			# Line.record / ops.pack_lines); the strip height is this pipeline's
			desc = lines.copy()
			desc["ysteps"] = self.line_height
		else:
			lines = np.ascontiguousarray(lines, dtype=np.float64).reshape(-1, 3, 2)
			desc = ops.pack_line_arrays(lines[:, 0], lines[:, 1], lines[:, 2], ysteps=self.line_height)
		n = desc.shape[0]
		w_max = 4
		if cap is None or self.hand_off == "device":
			est = _estimate_xsteps(grid, cell, desc) if n else np.zeros(0, np.int64)
			if cap is None:
				cap = int(np.sum(self.line_height * est)) if n else 1
			w_max = (int(est.max()) + 3) // 4 * 4 if n else 4
		if w_cap is not None:
			w_max = w_cap
		on_device = self.hand_off == "device"
		as_bits = self.hand_off == "bits"

		pitch = W  # the source is only read with byte loads: no row padding, one contiguous copy
		dpitch = (Wd + 127) // 128 * 128
		al = lambda v: (v + 255) // 256 * 256  # noqa: E731
		# input block: page rows (pitched) | grid | line descriptors
		o_grid = al(H * pitch)
		o_desc = o_grid + al(grid.nbytes)
		n_in = o_desc + al(max(desc.nbytes, 1))
		# output block: dewarped page (pitched) | strips | steps | pix offsets
		o_strips = al(Hd * dpitch) if self.want_dewarped else 0
		# device hand-off: the strips are not part of the download; bits: an eighth of them is
		o_steps = o_strips + (0 if on_device else al((cap + 7) // 8 + 4) if as_bits else al(cap))
		o_offs = o_steps + al(8 * max(n, 1))
		o_meta = o_offs + al(8 * (n + 1))  # widths | kept indices | n_kept (device hand-off)
		n_out = o_meta + (al(4 * (2 * max(n, 1) + 1)) if on_device else 0)

		if slot.done is not None:
			slot.done.synchronize()  # the slot's previous results must have been consumed
		slot.h_in = _Slot._grow(slot.h_in, n_in, pin_memory=True)
		slot.h_out = _Slot._grow(slot.h_out, n_out, pin_memory=True)
		slot.d_in = _Slot._grow(slot.d_in, n_in, device=self.device)
		slot.d_out = _Slot._grow(slot.d_out, n_out, device=self.device)
		slot.d_gray = _Slot._grow(slot.d_gray, cap, device=self.device)
		if on_device or as_bits:
			slot.d_strips = _Slot._grow(slot.d_strips, cap, device=self.device)
		if on_device:
			slot.d_batch = _Slot._grow(slot.d_batch, max(n, 1) * self.line_height * w_max, device=self.device)
			slot.d_slots = _Slot._grow(slot.d_slots, 4 * max(n, 1), device=self.device)
		slot.d_chunks = _Slot._grow(slot.d_chunks, 4 * (n + 2), device=self.device)
		if self.use_map:
			slot.d_map = _Slot._grow(slot.d_map, Hd * Wd * 8, device=self.device)

		hin = slot.h_in.numpy()
		pinned_src = _PINNED.get(page.ctypes.data)
		if pinned_src is not None and tuple(pinned_src.shape) != (H, W):
			pinned_src = None
		if pinned_src is None:
			_parallel_copy(hin, page.reshape(-1))
		hin[o_grid:o_grid + grid.nbytes] = grid.view(np.uint8).reshape(-1)
		if n:
			hin[o_desc:o_desc + desc.nbytes] = desc.view(np.uint8).reshape(-1)
		slot.keep = (page, pinned_src)  # the source must stay alive until the upload has run
		slot.job = job

		st = slot.stream
		with torch.cuda.stream(st):
			if pinned_src is None:
				slot.d_in[:n_in].copy_(slot.h_in[:n_in], non_blocking=True)
			else:
				slot.d_in[:H * W].copy_(pinned_src.reshape(-1), non_blocking=True)
				slot.d_in[o_grid:n_in].copy_(slot.h_in[o_grid:n_in], non_blocking=True)
			base_in, base_out = slot.d_in.data_ptr(), slot.d_out.data_ptr()
			s = st.cuda_stream
			p_map = slot.d_map.data_ptr() if self.use_map else 0
			if self.want_dewarped:
				_lib.call("og_remap_page", base_in, H, W, pitch, base_in + o_grid, gh, gw, cell,
					base_out, dpitch, p_map, 1.0, 1.0, ops.OG_INTER_LINEAR, s)
			elif n:
				_lib.call("og_expand_map", base_in + o_grid, gh, gw, cell, p_map, s)
			if n:
				_lib.call("og_line_steps", base_in + o_grid, gh, gw, cell, base_in + o_desc, n,
					base_out + o_steps, base_out + o_offs, slot.d_chunks.data_ptr(), s)
				mode, window = ops._binarize_args(self.binarize)
				scratch = slot.d_gray.data_ptr() if mode == ops.OG_BIN_SAUVOLA else 0
				p_strips = slot.d_strips.data_ptr() if (on_device or as_bits) else base_out + o_strips
				_lib.call("og_extract_lines", base_in, H, W, pitch, base_in + o_grid, gh, gw, cell, p_map,
					base_in + o_desc, n, base_out + o_steps, base_out + o_offs, slot.d_chunks.data_ptr(),
					p_strips, cap, 0, mode, window, scratch, 0, s)
				if as_bits:
					_lib.call("og_pack_bits", p_strips, cap, base_out + o_strips, s)
				if on_device:
					p_meta = base_out + o_meta
					_lib.call("og_pack_strips", p_strips, base_out + o_steps, base_out + o_offs, n, cap,
						self.min_size[0], self.min_size[1], self.pad_value, self.line_height, w_max,
						slot.d_batch.data_ptr(), slot.d_slots.data_ptr(), p_meta, p_meta + 4 * max(n, 1), p_meta + 8 * max(n, 1), s)
			slot.h_out[:n_out].copy_(slot.d_out[:n_out], non_blocking=True)
			slot.done = torch.cuda.Event()
			slot.done.record(st)
		slot.meta = (Hd, Wd, dpitch, n, cap, o_strips, o_steps, o_offs)
		slot.meta_dev = (o_meta, w_max)
		slot.bytes = (n_in, n_out)

	def _collect(self, slot):
		slot.done.synchronize()
		Hd, Wd, dpitch, n, cap, o_strips, o_steps, o_offs = slot.meta
		hout = slot.h_out.numpy()
		dewarped = hout[:Hd * dpitch].reshape(Hd, dpitch)[:, :Wd] if self.want_dewarped else None
		steps = hout[o_steps:o_steps + 8 * n].view(np.int32).reshape(n, 2)
		offs = hout[o_offs:o_offs + 8 * (n + 1)].view(np.int64)
		if n and int(offs[n]) > cap:
			# the host-side estimate of the strip widths fell short (it is not bit-exact): run the
			# page again with the size the device computed
			self._submit(slot, slot.job, cap=int(offs[n]))
			return self._collect(slot)
		if n and (steps <= 0).any():
			raise ValueError("line %d has a degenerate or non-finite extent" % int(np.argmax((steps <= 0).any(axis=1))))
		if self.hand_off == "device":
			o_meta, w_max = slot.meta_dev
			m = max(n, 1)
			meta = hout[o_meta:o_meta + 4 * (2 * m + 1)].view(np.int32)
			keep = (steps[:, 0] >= self.min_size[0]) & (steps[:, 1] >= self.min_size[1]) if n else np.zeros(0, bool)
			n_kept = int(meta[2 * m]) if n else 0
			if n_kept != int(keep.sum()):
				# a strip was wider than the batch sized from the host-side estimate: again, with the exact width
				self._submit(slot, slot.job, cap=max(cap, int(offs[n])), w_cap=(int(steps[keep, 0].max()) + 3) // 4 * 4)
				return self._collect(slot)
			batch = slot.d_batch[:n_kept * self.line_height * w_max].view(n_kept, self.line_height, w_max)
			return PageResult(dewarped, DeviceBatch(batch, meta[:n_kept].copy(), meta[m:m + n_kept].copy()), steps)
		if self.hand_off == "bits":
			return PageResult(dewarped, PackedStripList(hout[o_strips:o_strips + (cap + 7) // 8], offs, steps), steps)
		flat = hout[o_strips:o_strips + cap]
		return PageResult(dewarped, StripList(flat, offs, steps), steps)

	def run(self, jobs):
		"""Generator over results, in job order; `depth` pages are in flight."""
		pending = collections.deque()
		k = 0
		for job in jobs:
			slot = self.slots[k % len(self.slots)]
			if len(pending) == len(self.slots):
				yield self._collect(pending.popleft())
			self._submit(slot, job)
			pending.append(slot)
			k += 1
		while pending:
			yield self._collect(pending.popleft())

	@property
	def bytes_per_page(self):
		"""(host->device, device->host) bytes of the most recently submitted page."""
		return self.slots[0].bytes
