import sys, os, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from origami_b200 import ops, synth
dev = torch.device("cuda:0")
page = synth.make_page(4800, 7200, seed=0, kind="text")
big = torch.empty((84 << 20,), dtype=torch.uint8, device=dev)
def run(with_d2h, n=12):
	ts = []
	for i in range(n):
		torch.cuda.synchronize()
		t0 = time.perf_counter()
		d = ops.to_device_image(page, dev)
		torch.cuda.synchronize()
		ts.append(time.perf_counter() - t0)
		if with_d2h:
			h = ops.to_host(big)
			_ = int(h[-1])
	return np.median(ts) * 1e3
print("upload alone           %.2f ms" % run(False))
print("upload between D2H     %.2f ms" % run(True))
print("upload alone again     %.2f ms" % run(False))
