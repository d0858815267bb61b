"""Data-parallel plumbing for the codec: one process per GPU, launched by
torchrun; images are independent units (one string per image, `entropy_models.py:196-197`), so
the batch is split contiguously across ranks and there is NO collective on the coding path.
torch.distributed is used only around it: broadcasting externally loaded weights once,
gathering the host strings, and the max-over-ranks reduction of a timing.

Everything here is backend-agnostic (`nccl` on the GPUs, `gloo` in the CPU tests).
"""
import torch
import torch.distributed as dist

__all__ = ["shard_bounds", "local_shard", "gather_strings", "broadcast_tensors",
           "max_over_ranks", "world_info"]


def world_info():
  """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
  if dist.is_available() and dist.is_initialized():
    return dist.get_rank(), dist.get_world_size()
  return 0, 1


def shard_bounds(total, world):
  """Contiguous, balanced split of `total` units over `world` ranks: the first total % world
  ranks get one more.  Returns [(start, stop)] * world; the union is range(total)."""
  total, world = int(total), int(world)
  if world < 1 or total < 0:
    raise ValueError("shard_bounds: world must be >= 1 and total >= 0")
  base, extra = divmod(total, world)
  out, start = [], 0
  for r in range(world):
    n = base + (1 if r < extra else 0)
    out.append((start, start + n))
    start += n
  return out


def local_shard(total, rank=None, world=None):
  """(start, stop) of this rank's images."""
  if rank is None or world is None:
    rank, world = world_info()
  return shard_bounds(total, world)[rank]


def gather_strings(local_strings, dst=0, group=None):
  """Every rank passes the `bytes` strings of its shard (in image order); rank `dst` gets the
  strings of the whole batch in image order, the others get None.  Host-side: the strings are
  tiny next to the images and the reference hands them to the caller as host `bytes` too
  (`bmshj2018.py:339-347`)."""
  rank, world = world_info()
  local_strings = list(local_strings)
  if world == 1:
    return local_strings
  gathered = [None] * world if rank == dst else None
  dist.gather_object(local_strings, gathered, dst=dst, group=group)
  if rank != dst:
    return None
  return [s for part in gathered for s in part]


def broadcast_tensors(tensors, src=0, group=None):
  """In-place broadcast of a list of tensors (weights / CDF tables loaded on rank `src`).
  Ranks that build their parameters deterministically from the seed do not need it."""
  _, world = world_info()
  if world == 1:
    return tensors
  for t in tensors:
    dist.broadcast(t, src=src, group=group)
  return tensors


def max_over_ranks(value, device=None, group=None):
  """max of a Python float over all ranks (a step is as slow as its slowest rank)."""
  _, world = world_info()
  if world == 1:
    return float(value)
  t = torch.tensor([float(value)], dtype=torch.float64,
                   device=device if device is not None else "cpu")
  dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
  return float(t.item())
