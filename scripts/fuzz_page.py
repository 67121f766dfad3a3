"""Random page sizes, cells and warps: expand + remap (LINEAR, with and without the map, NEAREST at another
resolution), page Otsu / Sauvola and the strips of a few lines against the NumPy oracle (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import exact
from origami_b200 import ops, synth

dev = torch.device("cuda:0")
bad = 0
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 20):
	rng = np.random.default_rng(100 + seed)
	cell = int(rng.choice([7, 16, 25, 32]))
	W, H = int(rng.integers(90, 900)), int(rng.integers(90, 1100))
	page = synth.make_page(W, H, seed=seed, kind="noise" if seed % 2 else "text")
	grid = synth.make_grid(W, H, cell, seed=seed)
	gh, gw = grid.shape[:2]
	g = torch.from_numpy(grid).to(dev)
	src = ops.to_device_image(page, dev)
	dense = exact.expand_map(grid, cell)
	dst, mp = ops.remap_page(src, g, cell, map_out=torch.empty((gh * cell, gw * cell, 2), dtype=torch.float32, device=dev))
	res = {"map": np.array_equal(mp.cpu().numpy().view(np.uint32), dense.view(np.uint32)),
		"remap": np.array_equal(dst.cpu().numpy(), exact.remap_linear(page, dense))}
	dst2, _ = ops.remap_page(src, g, cell)
	res["remap_nomap"] = torch.equal(dst, dst2)
	lab = rng.integers(0, 5, size=(H // 2, W // 2), dtype=np.uint8)
	got, _ = ops.remap_page(ops.to_device_image(lab, dev), g, cell, scale=((W // 2) / W, (H // 2) / H), interpolation=ops.OG_INTER_NEAREST)
	res["nearest"] = np.array_equal(got.cpu().numpy(), exact.dewarp_image(lab, (W, H), grid, cell, interpolation=exact.INTER_NEAREST))
	win = int(rng.choice([3, 7, 15, 21]))
	res["sauvola"] = np.array_equal(ops.sauvola(src, win).cpu().numpy(), exact.binarize(page, ("sauvola", win)))
	res["otsu"] = np.array_equal(ops.otsu(src)[0].cpu().numpy(), exact.binarize(page, ("otsu",)))
	lines = synth.make_lines(gw * cell, gh * cell, 6, seed=seed, dpi=300)
	plan = ops.line_steps(g, cell, ops.lines_to_device(ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=24) for l in lines]), dev), len(lines))
	offs = plan.pix_offsets.cpu().numpy()
	out = torch.zeros((int(offs[-1]),), dtype=torch.uint8, device=dev)
	ops.extract_lines(src, g, cell, plan, out, dense_map=mp)
	o = out.cpu().numpy()
	ok = True
	for i, l in enumerate(lines):
		want = exact.dewarped_line_image(page, dense, l["p"], l["right"], l["up"], target_height=24)
		ok = ok and np.array_equal(o[offs[i]:offs[i + 1]].reshape(want.shape), want)
	res["strips"] = ok
	bad += 0 if all(res.values()) else 1
	print("seed %2d  %dx%d cell %2d window %2d  %s" % (seed, W, H, cell, win, res))
print("FAILED" if bad else "all equal")
