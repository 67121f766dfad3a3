"""Copy-only ceiling of the end-to-end page pipeline: every rank moves what PagePipeline moves per
page (H2D 35.2 MB from pinned memory, D2H 84.3 MB into pinned memory, three pages in flight on
three streams) with NO kernels.  If this takes as long per page as the pipeline's e2e number, the
host's DMA path is the limit, not the code.  Launch like bench.py (python, or torchrun for N > 1).
Prints one JSON line per run on rank 0."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from origami_b200 import shard

def main():
	rank, local_rank, world = shard.init()
	torch.cuda.set_device(local_rank)
	dev = torch.device("cuda", local_rank)
	bind = os.environ.get("OG_BIND_CORES")  # e.g. "per-rank": give every rank its own block of cores
	if bind:
		n = os.cpu_count() or 1
		per = max(1, n // world)
		os.sched_setaffinity(0, set(range(rank * per, min(n, (rank + 1) * per))))
	h2d_b, d2h_b, depth, pages = 35_234_304, 84_300_000, 3, int(os.environ.get("PAGES", 60))
	hin = [torch.empty(h2d_b, dtype=torch.uint8).pin_memory() for _ in range(depth)]
	hout = [torch.empty(d2h_b, dtype=torch.uint8).pin_memory() for _ in range(depth)]
	din = [torch.empty(h2d_b, dtype=torch.uint8, device=dev) for _ in range(depth)]
	dout = [torch.empty(d2h_b, dtype=torch.uint8, device=dev) for _ in range(depth)]
	for t in hin + hout:
		t.fill_(1)  # first touch after the affinity is set
	streams = [torch.cuda.Stream(device=dev) for _ in range(depth)]
	def run(n):
		for k in range(n):
			s = k % depth
			with torch.cuda.stream(streams[s]):
				din[s].copy_(hin[s], non_blocking=True)
				hout[s].copy_(dout[s], non_blocking=True)
		torch.cuda.synchronize()
	run(6)
	shard.barrier()
	t0 = time.perf_counter()
	run(pages)
	dt = shard.max_over_ranks((time.perf_counter() - t0) / pages, dev)
	if rank == 0:
		print(json.dumps({"n_gpus": world, "bind": bind or "default", "ms_per_page_and_gpu": 1e3 * dt,
			"aggregate_GBps": world * (h2d_b + d2h_b) / dt / 1e9, "h2d_bytes": h2d_b, "d2h_bytes": d2h_b}))
	shard.finalize()

if __name__ == "__main__":
	main()
