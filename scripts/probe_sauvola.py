"""Page Sauvola / Otsu alone (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops, synth
from probe_perf import timeit
dev = torch.device("cuda:0")
page = ops.to_device_image(synth.make_page(4800, 7200, seed=2, kind="text"), dev)
dst = ops.alloc_image(7200, 4800, dev)
if "--kernel-only" in sys.argv:
	for _ in range(3): ops.sauvola(page, 15, dst=dst)
	torch.cuda.synchronize()
else:
	ops.sauvola(page, 15, dst=dst); torch.cuda.synchronize()
	print("checksum", int(dst.to(torch.int64).sum().item()))
	for w in (15, 31):
		print("sauvola", w, timeit(lambda: ops.sauvola(page, w, dst=dst)))
	print("otsu", timeit(lambda: ops.otsu(page, dst=dst)))
