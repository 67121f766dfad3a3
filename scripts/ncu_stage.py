import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops, synth
W, H, cell = 4800, 7200, 25
dev = torch.device("cuda:0")
page = synth.make_page(W, H, seed=0, kind="noise")
grid = synth.make_grid(W, H, cell, seed=0)
g = torch.from_numpy(grid).to(dev); src = ops.to_device_image(page, dev)
mp = torch.empty((grid.shape[0]*cell, grid.shape[1]*cell, 2), dtype=torch.float32, device=dev)
for _ in range(3):
	ops.remap_page(src, g, cell, map_out=mp, interpolation=ops.OG_INTER_LINEAR | ops.OG_FLAG_STAGE_SMEM)
torch.cuda.synchronize()
