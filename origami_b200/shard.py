"""
Page sharding across GPUs: pages are independent, so rank r of N
takes pages r, r+N, r+2N, ... and no data-path collective exists.  torch.distributed is
used only for the start/end barriers and the max-over-ranks of the elapsed time.
"""

import os

import torch
import torch.distributed as dist


def env_rank():
	return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def pages_of_rank(n_pages, rank, world_size):
	"""Static page-index shard: page i -> rank i mod N (the reference's analogue is
	Processor's multiprocessing pool over page chunks, origami/batch/core/processor.py:470-500)."""
	return list(range(rank, n_pages, world_size))


def init(backend=None):
	"""Initialises the default process group from the torchrun environment (no-op for one rank)."""
	rank, local_rank, world = env_rank()
	if world > 1 and not dist.is_initialized():
		if backend is None:
			backend = "nccl" if torch.cuda.is_available() else "gloo"
		os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
		os.environ.setdefault("MASTER_PORT", "29500")
		kwargs = {}
		if backend == "nccl":
			kwargs["device_id"] = torch.device("cuda", local_rank)
		dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
	return rank, local_rank, world


def barrier():
	if dist.is_initialized():
		dist.barrier()


def max_over_ranks(value, device=None):
	"""max of a Python float over all ranks (the timing reduction of the bench contract)."""
	if not dist.is_initialized():
		return float(value)
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
	t = torch.tensor([float(value)], dtype=torch.float64, device=device)
	dist.all_reduce(t, op=dist.ReduceOp.MAX)
	return float(t.item())


def sum_over_ranks(value, device=None):
	if not dist.is_initialized():
		return float(value)
	if device is None:
		device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
	t = torch.tensor([float(value)], dtype=torch.float64, device=device)
	dist.all_reduce(t, op=dist.ReduceOp.SUM)
	return float(t.item())


def finalize():
	if dist.is_initialized():
		dist.destroy_process_group()
