"""Quick device-side timing of the individual kernels at BASELINE sizes (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth

def timeit(fn, iters=20, warm=3):
	"""Median / best device time of fn(), launches back to back.  A few milliseconds of a spinning
	kernel are queued first, so that all timed launches are already in the stream when the GPU
	reaches them: without that a kernel shorter than the host's launch cost (~40 us through ctypes)
	runs with idle gaps around it, which flatters bandwidth-bound kernels (the write-back of the
	previous launch drains in the gap) -- and a compute-heavy filler would skew the clocks."""
	for _ in range(warm): fn()
	torch.cuda.synchronize()
	ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
	torch.cuda._sleep(8_000_000)
	for a, b in ev:
		a.record(); fn(); b.record()
	torch.cuda.synchronize()
	ts = sorted(a.elapsed_time(b) for a, b in ev)
	return ts[len(ts) // 2], ts[0]

def main():
	W, H, cell = (int(os.environ.get('PW', 4800)), int(os.environ.get('PH', 7200)), 25)
	n_lines = int(os.environ.get('PLINES', 2000))
	dev = torch.device("cuda:0")
	page = synth.make_page(W, H, seed=0, kind="text")
	grid = synth.make_grid(W, H, cell, seed=0)
	gh, gw = grid.shape[:2]
	Hd, Wd = gh * cell, gw * cell
	lines = synth.make_lines(Wd, Hd, n_lines, seed=0, dpi=600)
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	dst = ops.alloc_image(Hd, Wd, dev)
	mp = torch.empty((Hd, Wd, 2), dtype=torch.float32, device=dev)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines])
	ld = ops.lines_to_device(la, dev)
	plan = ops.line_steps(g, cell, ld, n_lines)
	total = int(plan.pix_offsets[-1].item())
	strips = torch.empty((total,), dtype=torch.uint8, device=dev)
	strips2 = torch.empty_like(strips)
	bbox = torch.empty((n_lines, 4), dtype=torch.int32, device=dev)
	print("page %dx%d dewarped %dx%d lines %d strip px %.1f MP" % (W, H, Wd, Hd, n_lines, total / 1e6))
	res = {}
	res["expand_map"] = timeit(lambda: ops.expand_map(g, cell, out=mp)), 8 * Hd * Wd
	res["remap+map"] = timeit(lambda: ops.remap_page(src, g, cell, dst=dst, map_out=mp)), 9 * Hd * Wd + H * W
	res["remap only"] = timeit(lambda: ops.remap_page(src, g, cell, dst=dst)), Hd * Wd + H * W
	res["line_steps"] = timeit(lambda: ops.line_steps(g, cell, ld, n_lines)), 0
	res["extract"] = timeit(lambda: ops.extract_lines(src, g, cell, plan, strips, bbox)), 2 * total
	res["extract nobbox"] = timeit(lambda: ops.extract_lines(src, g, cell, plan, strips)), 2 * total
	res["extract map"] = timeit(lambda: ops.extract_lines(src, g, cell, plan, strips, bbox, dense_map=mp)), 2 * total
	res["extract map nobbox"] = timeit(lambda: ops.extract_lines(src, g, cell, plan, strips, dense_map=mp)), 2 * total
	res["otsu strips"] = timeit(lambda: ops.binarize_strips(strips, plan, ops.OG_BIN_OTSU, out=strips2)), 2 * total
	res["sauvola strips"] = timeit(lambda: ops.binarize_strips(strips, plan, ops.OG_BIN_SAUVOLA, 15, out=strips2)), 2 * total
	bin_dst = ops.alloc_image(H, W, dev)
	res["sauvola page"] = timeit(lambda: ops.sauvola(src, 15, dst=bin_dst)), 2 * H * W
	res["otsu page"] = timeit(lambda: ops.otsu(src, dst=bin_dst)), 3 * H * W
	probs = torch.rand((5, 2400, 1280, 4), device=dev)
	res["argmax M5"] = timeit(lambda: ops.argmax_labels(probs, [1, 2, 3])), (4 * 5 * 4 + 4) * 2400 * 1280
	la2 = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], xres=0.5, yres=0.5) for l in lines])
	ld2 = ops.lines_to_device(la2, dev)
	plan2 = ops.line_steps(g, cell, ld2, n_lines)
	labels = ops.to_device_image(synth.make_labels(1280, 2400), dev)
	res["sample labels"] = timeit(lambda: ops.sample_labels(labels, (H, W), g, cell, plan2, 4)), 0
	res["sample labels map"] = timeit(lambda: ops.sample_labels(labels, (H, W), g, cell, plan2, 4, dense_map=mp)), 0
	for k, ((med, best), nbytes) in res.items():
		print("%-16s median %8.3f ms  best %8.3f ms  %8.1f GB/s" % (k, med, best, nbytes / best / 1e6))

if __name__ == "__main__":
	main()
