"""Time only the line kernels (development aid)."""
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
	ld = ops.lines_to_device(la, dev)
	plan = ops.line_steps(g, cell, ld, n_lines)
	total = int(plan.pix_offsets[-1].item())
	strips = torch.empty((total,), dtype=torch.uint8, device=dev)
	bbox = torch.empty((n_lines, 4), dtype=torch.int32, device=dev)
	for name, fn in [
		("extract staged", lambda: ops.extract_lines(src, g, cell, plan, strips)),
		("extract map", lambda: ops.extract_lines(src, g, cell, plan, strips, dense_map=mp)),
		("extract map bbox", lambda: ops.extract_lines(src, g, cell, plan, strips, bbox, dense_map=mp)),
		("line_steps", lambda: ops.line_steps(g, cell, ld, n_lines)),
		("remap+map", lambda: ops.remap_page(src, g, cell, map_out=mp)),
		("remap+map staged", lambda: ops.remap_page(src, g, cell, map_out=mp, interpolation=ops.OG_INTER_LINEAR | ops.OG_FLAG_STAGE_SMEM)),
	]:
		med, best = timeit(fn)
		print("%-20s %s median %.3f ms best %.3f ms" % (name, os.environ.get("ORIGAMI_B200_LIB", "default")[-12:], med, best))

if __name__ == "__main__":
	main()
