"""Cross-check of the strip engines and binarizers over random pages, grids and lines (development aid):
screened map-gather vs float64 map-gather vs line-owning (no map) must agree byte for byte; Otsu and
Sauvola through both routes; warp Sauvola vs the tile kernel on the page."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from origami_b200 import ops, synth

dev = torch.device("cuda:0")
bad = 0
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
	rng = np.random.default_rng(seed)
	cell = int(rng.choice([16, 20, 25, 32]))
	W, H = int(rng.integers(1200, 4800)), int(rng.integers(1500, 7200))
	n = int(rng.integers(50, 1500))
	page = synth.make_page(W, H, seed=seed, kind="text" if seed % 3 else "noise")
	grid = synth.make_grid(W, H, cell, seed=seed)
	gh, gw = grid.shape[:2]
	lines = synth.make_lines(gw * cell, gh * cell, n, seed=seed, dpi=int(rng.choice([300, 600])))
	hgt = int(rng.choice([32, 48, 56]))
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	mp = ops.expand_map(g, cell)
	plan = ops.line_steps(g, cell, ops.lines_to_device(ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=hgt) for l in lines]), dev), n)
	total = int(plan.pix_offsets[-1].item())
	out = [torch.zeros((total,), dtype=torch.uint8, device=dev) for _ in range(3)]
	ops.extract_lines(src, g, cell, plan, out[0])                                   # line-owning, no map
	ops.extract_lines(src, g, cell, plan, out[1], dense_map=mp)                     # screened
	bb = torch.zeros((n, 4), dtype=torch.int32, device=dev)
	ops.extract_lines(src, g, cell, plan, out[2], dense_map=mp, bbox=bb)            # float64 map-gather
	ok = torch.equal(out[0], out[1]) and torch.equal(out[1], out[2])
	res = {}
	for spec in (("otsu",), ("sauvola", 15)):
		a, b = torch.zeros_like(out[0]), torch.zeros_like(out[0])
		gray = torch.zeros_like(out[0])
		ops.extract_lines(src, g, cell, plan, a, binarize=spec, gray_out=gray)
		ops.extract_lines(src, g, cell, plan, b, dense_map=mp, binarize=spec, gray_out=gray)
		res[spec[0]] = torch.equal(a, b)
	os.environ.pop("OG_SAUVOLA_LEGACY", None)
	s1 = ops.sauvola(src, 15)
	os.environ["OG_SAUVOLA_LEGACY"] = "1"
	s2 = ops.sauvola(src, 15)
	os.environ.pop("OG_SAUVOLA_LEGACY", None)
	res["sauvola_page"] = torch.equal(s1, s2)
	good = ok and all(res.values())
	bad += 0 if good else 1
	print("seed %2d  %dx%d cell %d  %4d lines h=%d  %.1f MP strips  engines %s  %s" % (seed, W, H, cell, n, hgt, total / 1e6, ok, res))
print("FAILED" if bad else "all equal")
