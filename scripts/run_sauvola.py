"""A few launches of page Sauvola-15 on the bench page (target for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops, synth
W, H = 4800, 7200
dev = torch.device("cuda:0")
page = synth.make_page(W, H, seed=0, kind="text")
src = ops.to_device_image(page, dev)
dst = ops.alloc_image(H, W, dev)
for _ in range(4):
	ops.sauvola(src, 15, dst=dst)
torch.cuda.synchronize()
print("ok")
