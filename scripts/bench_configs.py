"""Device-side timings of the other BASELINE.json configurations (parity cases, not bench lines):
configs[0] 2400x3600 page, configs[3] line-extraction stress, configs[4] segmentation post-process.
Writes a small report to stdout; CUDA events, median of 20 after 3 warm-ups."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth
from probe_perf import timeit


def page_path(W, H, n_lines, dpi, dev, cell=25):
	page = synth.make_page(W, H, seed=1, kind="text", dpi=dpi)
	grid = synth.make_grid(W, H, cell, seed=1)
	gh, gw = grid.shape[:2]
	Hd, Wd = gh * cell, gw * cell
	lines = synth.make_lines(Wd, Hd, n_lines, seed=1, dpi=dpi)
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	dst = ops.alloc_image(Hd, Wd, dev)
	mp = torch.empty((Hd, Wd, 2), dtype=torch.float32, device=dev)
	ld = ops.lines_to_device(ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines]), dev)
	plan = ops.line_steps(g, cell, ld, n_lines)
	total = int(plan.pix_offsets[-1].item())
	gray = torch.empty((total,), dtype=torch.uint8, device=dev)
	out = torch.empty_like(gray)

	def step(binarize):
		ops.remap_page(src, g, cell, dst=dst, map_out=mp)
		p = ops.line_steps(g, cell, ld, n_lines)
		ops.extract_lines(src, g, cell, p, gray, dense_map=mp)
		if binarize == "otsu":
			ops.binarize_strips(gray, p, ops.OG_BIN_OTSU, out=out)
		elif binarize == "sauvola":
			ops.binarize_strips(gray, p, ops.OG_BIN_SAUVOLA, 15, out=out)
	res = {}
	for b in ("none", "otsu", "sauvola"):
		med, best = timeit(lambda: step(b))
		res[b] = dict(ms=med, page_MPs=Hd * Wd / med / 1e3, strip_MPs=total / med / 1e3)
	return dict(page=[W, H], dewarped=[Wd, Hd], lines=n_lines, strip_px=total, steps=res)


def seg_post(dev):
	out = {}
	for M in (1, 5):
		probs = torch.rand((M, 2400, 1280, 4), device=dev)
		med, _ = timeit(lambda: ops.argmax_labels(probs, [1, 2, 3]))
		out["argmax_M%d_ms" % M] = med
		out["argmax_M%d_GBs" % M] = (4 * M * 4 + 4) * 2400 * 1280 / med / 1e6
	page = ops.to_device_image(synth.make_page(4800, 7200, seed=2, kind="text"), dev)
	dst = ops.alloc_image(7200, 4800, dev)
	med, _ = timeit(lambda: ops.sauvola(page, 15, dst=dst))
	out["sauvola15_page_ms"] = med
	med, _ = timeit(lambda: ops.otsu(page, dst=dst))
	out["otsu_page_ms"] = med
	out["pages_per_s_argmaxM5_plus_sauvola"] = 1e3 / (out["argmax_M5_ms"] + out["sauvola15_page_ms"])
	return out


def main():
	dev = torch.device("cuda:0")
	rep = {
		"configs[0] 2400x3600 page, 1200 lines (300 dpi)": page_path(2400, 3600, 1200, 300, dev),
		"configs[1]/[3] 4800x7200 page, 2000 lines (600 dpi)": page_path(4800, 7200, 2000, 600, dev),
		"configs[4] segmentation post-process (label map 1280x2400, page 4800x7200)": seg_post(dev),
	}
	print(json.dumps(rep, indent=1))


if __name__ == "__main__":
	main()
