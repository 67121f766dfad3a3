"""Worker for tests/test_host_logic.py::test_sharding_two_ranks_gloo (run under torchrun)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from origami_b200 import shard  # noqa: E402


def main():
	rank, local_rank, world = shard.init(backend="gloo")
	pages = shard.pages_of_rank(7, rank, world)
	shard.barrier()
	# each rank reports a different "elapsed time"; the job time is the max over ranks
	t = shard.max_over_ranks(10.0 + rank)
	total = shard.sum_over_ranks(len(pages))
	out = dict(rank=rank, world=world, pages=pages, tmax=t, total=total)
	with open(os.path.join(sys.argv[1], "rank%d.json" % rank), "w") as f:
		json.dump(out, f)
	shard.finalize()


if __name__ == "__main__":
	main()
