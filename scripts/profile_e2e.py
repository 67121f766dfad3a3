"""cProfile of the end-to-end (host buffers) path of bench.py (development aid)."""
import cProfile, pstats, sys, os, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

args = types.SimpleNamespace(lines=2000, binarize="otsu", page_sets=1)
inputs = [bench.make_inputs(args, 0)]
dev = torch.device("cuda:0")
bench.run_e2e(args, inputs, dev, 1)  # warm
pr = cProfile.Profile()
pr.enable()
s, h2d, d2h = bench.run_e2e(args, inputs, dev, 3)
pr.disable()
print("e2e ms/step", s * 1e3)
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
