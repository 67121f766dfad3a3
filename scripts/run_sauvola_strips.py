"""A few launches of strip Sauvola-15 on the bench page's strips (target for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops, synth
W, H, cell, n = 4800, 7200, 25, 2000
dev = torch.device("cuda:0")
page = synth.make_page(W, H, seed=0, kind="text")
grid = synth.make_grid(W, H, cell, seed=0)
gh, gw = grid.shape[:2]
lines = synth.make_lines(gw * cell, gh * cell, n, seed=0, dpi=600)
g = torch.from_numpy(grid).to(dev)
src = ops.to_device_image(page, dev)
mp = ops.expand_map(g, cell)
plan = ops.line_steps(g, cell, ops.lines_to_device(ops.pack_lines([dict(p=l["p"], right=l["right"], up=l["up"], ysteps=48) for l in lines]), dev), n)
total = int(plan.pix_offsets[-1].item())
gray = torch.empty((total,), dtype=torch.uint8, device=dev)
out = torch.empty_like(gray)
ops.extract_lines(src, g, cell, plan, gray, dense_map=mp)
for _ in range(3):
	ops.binarize_strips(gray, plan, ops.OG_BIN_SAUVOLA, 15, out=out)
torch.cuda.synchronize()
print("ok", total)
