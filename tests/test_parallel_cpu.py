"""Host-side data-parallel logic on CPU: world_size-2 `gloo` processes.

The coding path has no collective; what is tested is what surrounds it — the contiguous batch
split, the gather of the per-image strings in image order, the one-off broadcast of loaded
parameters and the max-over-ranks timing reduction bench.py uses.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from compression_b200 import parallel


def _free_port():
  s = socket.socket()
  s.bind(("127.0.0.1", 0))
  port = s.getsockname()[1]
  s.close()
  return port


def _fake_string(image):
  """Deterministic stand-in for one image's bitstring (length and content depend on it)."""
  rng = np.random.default_rng(1000 + image)
  return rng.integers(0, 256, size=10 + 3 * image, dtype=np.uint8).tobytes()


def _worker(rank, world, port, total, q):
  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  dist.init_process_group("gloo", rank=rank, world_size=world)
  try:
    assert parallel.world_info() == (rank, world)
    lo, hi = parallel.local_shard(total)
    mine = [_fake_string(i) for i in range(lo, hi)]
    got = parallel.gather_strings(mine, dst=0)
    w = torch.full((5,), float(rank + 7))
    parallel.broadcast_tensors([w], src=0)
    slowest = parallel.max_over_ranks(10.0 + rank)
    q.put((rank, (lo, hi), got, w.tolist(), slowest))
  finally:
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 8, 1])
def test_two_rank_shard_gather_broadcast_max(total):
  world, port = 2, _free_port()
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
  for p in procs:
    p.start()
  out = sorted([q.get(timeout=120) for _ in procs])
  for p in procs:
    p.join(timeout=60)
    assert p.exitcode == 0
  bounds = [o[1] for o in out]
  assert bounds == parallel.shard_bounds(total, world)
  assert bounds[0][0] == 0 and bounds[-1][1] == total and bounds[0][1] == bounds[1][0]
  # rank 0 holds every image's string, in image order; the other rank holds nothing
  assert out[0][2] == [_fake_string(i) for i in range(total)]
  assert out[1][2] is None
  assert out[0][3] == [7.0] * 5 and out[1][3] == [7.0] * 5       # broadcast from rank 0
  assert out[0][4] == 11.0 and out[1][4] == 11.0                 # max over ranks


def test_shard_bounds_properties():
  for total in (0, 1, 5, 64, 255, 256):
    for world in (1, 2, 3, 4, 8):
      b = parallel.shard_bounds(total, world)
      assert len(b) == world and b[0][0] == 0 and b[-1][1] == total
      sizes = [hi - lo for lo, hi in b]
      assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
      assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
  with pytest.raises(ValueError):
    parallel.shard_bounds(4, 0)


def test_single_process_defaults():
  assert parallel.world_info() == (0, 1)
  assert parallel.local_shard(10) == (0, 10)
  assert parallel.gather_strings([b"a", b"b"]) == [b"a", b"b"]
  assert parallel.max_over_ranks(3.5) == 3.5
