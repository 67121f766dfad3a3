"""A few launches of the fused page kernel on the bench page (target for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops, synth
W, H, cell = 4800, 7200, 25
dev = torch.device("cuda:0")
page = synth.make_page(W, H, seed=0, kind="text")
grid = synth.make_grid(W, H, cell, seed=0)
gh, gw = grid.shape[:2]
g = torch.from_numpy(grid).to(dev)
src = ops.to_device_image(page, dev)
dst = ops.alloc_image(gh * cell, gw * cell, dev)
mp = torch.empty((gh * cell, gw * cell, 2), dtype=torch.float32, device=dev)
for _ in range(4):
	ops.remap_page(src, g, cell, dst=dst, map_out=mp)
	if len(sys.argv) > 1:
		ops.remap_page(src, g, cell, dst=dst)
torch.cuda.synchronize()
print("ok")
