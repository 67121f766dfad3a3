"""Where the multi-page pipeline spends host time (development aid)."""
import cProfile, pstats, sys, os, time, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from origami_b200.batch.core.pipeline import PagePipeline, pinned_page

args = types.SimpleNamespace(lines=2000, binarize="otsu", page_sets=2)
inputs = [bench.make_inputs(args, k) for k in range(2)]
jobs = []
for page, g, lines in inputs:
	arr = np.array([[l["p"], l["right"], l["up"]] for l in lines], dtype=np.float64).reshape(-1, 3, 2)
	jobs.append((pinned_page(page), g, bench.CELL, arr))
# raw PCIe numbers
h = torch.empty((90 << 20,), dtype=torch.uint8, pin_memory=True); d = torch.empty_like(h, device="cuda")
for name, fn in (("H2D", lambda: d.copy_(h, non_blocking=True)), ("D2H", lambda: h.copy_(d, non_blocking=True))):
	fn(); torch.cuda.synchronize(); t = time.perf_counter(); fn(); torch.cuda.synchronize()
	print(name, "90 MiB: %.2f ms  %.1f GB/s" % ((time.perf_counter() - t) * 1e3, (90 << 20) / (time.perf_counter() - t) / 1e9))
# both directions at once, on two streams
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
h2 = torch.empty((35 << 20,), dtype=torch.uint8, pin_memory=True); d2 = torch.empty_like(h2, device="cuda")
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(10):
	with torch.cuda.stream(s1): h.copy_(d, non_blocking=True)
	with torch.cuda.stream(s2): d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 10
print("duplex: 90 MiB D2H + 35 MiB H2D per round: %.2f ms (D2H alone would be %.1f GB/s)" % (dt * 1e3, (90 << 20) / dt / 1e9))
for depth in (2, 3, 4):
	pipe = PagePipeline(48, "otsu", depth=depth)
	for _ in pipe.run(jobs): pass
	torch.cuda.synchronize(); t = time.perf_counter(); n = 24
	for _ in pipe.run(jobs[i % 2] for i in range(n)): pass
	torch.cuda.synchronize(); print("depth", depth, "ms/page %.2f" % ((time.perf_counter() - t) / n * 1e3))
pipe = PagePipeline(48, "otsu")
for _ in pipe.run(jobs): pass
pr = cProfile.Profile(); pr.enable()
for _ in pipe.run(jobs[i % 2] for i in range(12)): pass
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
