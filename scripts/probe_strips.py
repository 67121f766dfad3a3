"""Device-side timing and cross-check of the strip engines on the bench page (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth
from probe_perf import timeit

def main():
	W, H, cell = 4800, 7200, 25
	n_lines = int(os.environ.get('PLINES', 2000))
	dev = torch.device("cuda:0")
	page = synth.make_page(W, H, seed=0, kind=os.environ.get('PKIND', 'text'))
	grid = synth.make_grid(W, H, cell, seed=0)
	gh, gw = grid.shape[:2]
	lines = synth.make_lines(gw * cell, gh * cell, n_lines, seed=0, dpi=600)
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	mp = ops.expand_map(g, cell)
	plan = ops.line_steps(g, cell, ops.lines_to_device(ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines]), dev), n_lines)
	offs = plan.pix_offsets.cpu().numpy()
	steps = plan.steps.cpu().numpy()
	total = int(offs[-1])
	a = torch.zeros((total,), dtype=torch.uint8, device=dev)
	b = torch.zeros_like(a)
	c = torch.zeros_like(a)
	ops.extract_lines(src, g, cell, plan, a)
	ops.extract_lines(src, g, cell, plan, b, dense_map=mp, map_gather=True)
	an, bn = a.cpu().numpy(), b.cpu().numpy()
	bad = np.nonzero(an != bn)[0]
	print("line-owning vs screened: %d of %d pixels differ" % (bad.size, total))
	li = np.searchsorted(offs, bad, side="right") - 1
	for k in range(min(bad.size, 12)):
		i = li[k]
		row, col = divmod(int(bad[k] - offs[i]), int(steps[i, 0]))
		print("  line %d (xsteps %d) r=%d col %d (chunk col %d): %d vs %d" % (i, steps[i, 0], 47 - row, col, col % 64, an[bad[k]], bn[bad[k]]))
	ops.extract_lines(src, g, cell, plan, a, binarize=("otsu",))
	ops.extract_lines(src, g, cell, plan, b, dense_map=mp, binarize=("otsu",))
	print("otsu: equal =", torch.equal(a, b))
	for name, fn in [
		("line-owning none", lambda: ops.extract_lines(src, g, cell, plan, a)),
		("line-owning otsu", lambda: ops.extract_lines(src, g, cell, plan, a, binarize=("otsu",))),
		("screened none", lambda: ops.extract_lines(src, g, cell, plan, b, dense_map=mp, map_gather=True)),
		("screened + otsu", lambda: ops.extract_lines(src, g, cell, plan, b, dense_map=mp, binarize=("otsu",))),
	]:
		med, best = timeit(fn)
		print("%-18s median %8.3f ms  best %8.3f ms  (%.1f MP strips)" % (name, med, best, total / 1e6))

if __name__ == "__main__":
	main()
