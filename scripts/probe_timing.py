"""How the timing method changes what a bandwidth-bound kernel measures (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from origami_b200 import ops

dev = torch.device("cuda:0")
probs = torch.rand((5, 2400, 1280, 4), device=dev)
fn = lambda: ops.argmax_labels(probs, [1, 2, 3])
for _ in range(3): fn()
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
for a, b in ev:
	a.record(); fn(); b.record(); torch.cuda.synchronize()
print("isolated (sync after each):", sorted(a.elapsed_time(b) for a, b in ev)[10])
for a, b in ev:
	a.record(); fn(); b.record()
torch.cuda.synchronize()
print("queued by the host as fast as it can:", sorted(a.elapsed_time(b) for a, b in ev)[10])
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
busy = torch.empty((8192, 8192), dtype=torch.float16, device=dev).normal_()
torch.cuda.synchronize()
for _ in range(4): torch.mm(busy, busy)
a.record()
for _ in range(50): fn()
b.record(); torch.cuda.synchronize()
print("50 back to back behind queued work, per call:", a.elapsed_time(b) / 50)
x = torch.empty((64 << 20,), dtype=torch.float32, device=dev)
torch.cuda.synchronize()
a.record()
for _ in range(50): fn()
b.record(); torch.cuda.synchronize()
print("50 back to back, host-paced, per call:", a.elapsed_time(b) / 50)
torch.cuda.synchronize()
torch.cuda._sleep(20_000_000)
a.record()
for _ in range(50): fn()
b.record(); torch.cuda.synchronize()
print("50 back to back behind a sleep kernel, per call:", a.elapsed_time(b) / 50)
torch.cuda._sleep(20_000_000)
for p, q in ev:
	p.record(); fn(); q.record()
torch.cuda.synchronize()
print("per-call events behind a sleep kernel:", sorted(p.elapsed_time(q) for p, q in ev)[10])
