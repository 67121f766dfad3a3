"""Time strip Otsu / Sauvola on the bench workload (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth
from probe_perf import timeit

def main():
	W, H, cell, n_lines = 4800, 7200, 25, 2000
	dev = torch.device("cuda:0")
	page = synth.make_page(W, H, seed=0, kind="text")
	grid = synth.make_grid(W, H, cell, seed=0)
	gh, gw = grid.shape[:2]
	lines = synth.make_lines(gw * cell, gh * cell, n_lines, seed=0, dpi=600)
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	mp = ops.expand_map(g, cell)
	la = ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines])
	plan = ops.line_steps(g, cell, ops.lines_to_device(la, dev), n_lines)
	total = int(plan.pix_offsets[-1].item())
	gray = torch.zeros((total,), dtype=torch.uint8, device=dev)
	ops.extract_lines(src, g, cell, plan, gray, dense_map=mp)
	out = torch.empty_like(gray)
	th = torch.zeros((n_lines,), dtype=torch.int32, device=dev)
	ops.binarize_strips(gray, plan, ops.OG_BIN_OTSU, out=out, thresholds=th)
	torch.cuda.synchronize()
	print("checksum", int(out.to(torch.int64).sum().item()), int(th.to(torch.int64).sum().item()))
	for name, fn in [("otsu strips", lambda: ops.binarize_strips(gray, plan, ops.OG_BIN_OTSU, out=out)),
			("extract+otsu", lambda: (ops.extract_lines(src, g, cell, plan, gray, dense_map=mp), ops.binarize_strips(gray, plan, ops.OG_BIN_OTSU, out=out))),
			("line_steps", lambda: ops.line_steps(g, cell, plan.lines, n_lines))]:
		med, best = timeit(fn)
		print("%-16s %s median %.3f ms best %.3f ms" % (name, os.environ.get("ORIGAMI_B200_LIB", "default")[-14:], med, best))

if __name__ == "__main__":
	main()
