"""og_otsu_noalloc: the page-Otsu entry point for callers that cannot allocate device memory
(a stage plugin sharing its process with TensorFlow, which has already reserved the GPU; a worker
that only owns the buffers it was handed).  Its histogram scratch is a static device array inside
the library, shared by every call on the device, so the test makes two calls in a row on different
images -- the second would inherit the first one's counts if the entry point did not clear them --
and requires the same bytes and the same threshold as og_otsu with caller-provided scratch.
Kept in a file that pytest collects last."""

import numpy as np
import pytest
import torch

from oracle import exact
from origami_b200 import _lib, ops, synth

pytestmark = pytest.mark.gpu


def _otsu_noalloc(src):
	h, w = int(src.shape[0]), int(src.shape[1])
	dst = ops.alloc_image(h, w, src.device)
	thr = torch.full((1,), -1, dtype=torch.int32, device=src.device)
	_lib.call("og_otsu_noalloc", src.data_ptr(), h, w, src.stride(0), dst.data_ptr(), dst.stride(0),
		thr.data_ptr(), torch.cuda.current_stream().cuda_stream)
	return dst, thr


def test_noalloc_matches_og_otsu_on_two_calls_in_a_row(cuda):
	big = ops.to_device_image(synth.make_page(1500, 1100, seed=41, kind="text", dpi=300), cuda)
	small = ops.to_device_image(synth.make_page(333, 217, seed=42, kind="noise"), cuda)
	flat = ops.to_device_image(np.full((64, 96), 200, np.uint8), cuda)
	for src in (big, small, flat):  # back to back: the shared counters must start from zero each time
		want, want_t = ops.otsu(src)
		got, got_t = _otsu_noalloc(src)
		assert torch.equal(got, want)
		assert int(got_t.item()) == int(want_t.item())
		# and against the oracle itself, not only against the sibling entry point
		assert np.array_equal(got.cpu().numpy(), exact.binarize(src.cpu().numpy(), ("otsu",)))


def test_noalloc_rejects_images_beyond_its_limit(cuda):
	lib = _lib.load()
	buf = torch.zeros((16,), dtype=torch.uint8, device=cuda)
	rc = lib.og_otsu_noalloc(buf.data_ptr(), 40000, 40000, 40000, buf.data_ptr(), 40000, None, None)
	assert rc == -2
	assert b"limit" in lib.og_last_error()
