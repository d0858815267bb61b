"""Two ranks on two GPUs over NCCL: the data-parallel training step's gradient
all-reduce and the sharded codec.  Skipped on a box with fewer than two GPUs (the gloo
world-size-2 tests in test_parallel_cpu.py cover the host logic everywhere)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
  with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
    s.bind(("127.0.0.1", 0))
    return s.getsockname()[1]


def _need_two_gpus():
  if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
    pytest.skip("needs two GPUs")


def _train_worker(rank, world, port, q):
  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  os.environ.setdefault("NCCL_P2P_LEVEL", "NVL")
  os.environ.setdefault("NCCL_IB_DISABLE", "1")
  dev = torch.device("cuda", rank)
  torch.cuda.set_device(dev)
  dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
  try:
    from compression_b200.examples import bls2017
    from compression_b200.python.ops import train_ops
    model = bls2017.Bls2017Model(128, dev, seed=0)                # identical replicas
    trainer = bls2017.Bls2017Trainer(model, lmbda=0.01)
    g = torch.Generator(device=dev).manual_seed(100 + rank)       # per-rank crops and noise
    x = torch.rand((4, 64, 64, 3), generator=g, device=dev)
    y_shape = (4, 4, 4, 128)
    noise = torch.rand(y_shape, generator=g, device=dev) - .5
    before = trainer.main.param.clone()

    # the all-reduced gradient is the sum of the two local ones (two addends: exact)
    trainer._forward_backward(x, noise)
    local = trainer.main.grad.clone()
    both = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(both, local)
    scale = trainer.main.all_reduce()
    trainer.aux.all_reduce()
    assert scale == 1.0 / world
    assert torch.equal(trainer.main.grad, both[0] + both[1])
    assert float((both[0] - both[1]).abs().max()) > 0            # the ranks saw different data
    train_ops.adam_step(trainer.main, trainer.main_lr, scale)
    train_ops.adam_step(trainer.aux, trainer.aux_lr, scale)
    trainer._invalidate()
    model.entropy_bottleneck.update()

    # two more whole steps (CUDA graph replay + all-reduce + Adam): replicas stay identical
    for _ in range(2):
      stats = trainer.step(x, noise, graph=True)
    assert np.isfinite(stats["loss"])
    for bucket in (trainer.main, trainer.aux):
      mine = bucket.param.clone()
      theirs = [torch.empty_like(mine) for _ in range(world)]
      dist.all_gather(theirs, mine)
      assert torch.equal(theirs[0], theirs[1])
    assert float((trainer.main.param - before).abs().max()) > 0
    q.put((rank, "ok"))
  except Exception as exc:  # pylint: disable=broad-except
    q.put((rank, "%s: %s" % (type(exc).__name__, exc)))
  finally:
    dist.destroy_process_group()


def _codec_worker(rank, world, port, q):
  os.environ["MASTER_ADDR"] = "127.0.0.1"
  os.environ["MASTER_PORT"] = str(port)
  os.environ.setdefault("NCCL_P2P_LEVEL", "NVL")
  os.environ.setdefault("NCCL_IB_DISABLE", "1")
  dev = torch.device("cuda", rank)
  torch.cuda.set_device(dev)
  dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
  try:
    from compression_b200 import parallel
    from compression_b200.examples import bmshj2018
    model = bmshj2018.Bmshj2018Model(128, dev, seed=0)
    g = torch.Generator().manual_seed(7)
    images = torch.rand((6, 64, 96, 3), generator=g)              # the same batch on every rank
    lo, hi = parallel.local_shard(images.shape[0])
    mine = model.to_host(model.compress(images[lo:hi].to(dev)))
    strings = parallel.gather_strings(list(zip(mine.string, mine.side_string)))
    t = parallel.max_over_ranks(float(rank), device=dev)
    assert t == float(world - 1)
    if rank == 0:
      whole = model.to_host(model.compress(images.to(dev)))
      assert strings == list(zip(whole.string, whole.side_string))
    else:
      assert strings is None
    q.put((rank, "ok"))
  except Exception as exc:  # pylint: disable=broad-except
    q.put((rank, "%s: %s" % (type(exc).__name__, exc)))
  finally:
    dist.destroy_process_group()


def _run(worker):
  _need_two_gpus()
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = _free_port()
  procs = [ctx.Process(target=worker, args=(r, 2, port, q)) for r in range(2)]
  for p in procs:
    p.start()
  results = {}
  for _ in procs:
    rank, msg = q.get(timeout=600)
    results[rank] = msg
  for p in procs:
    p.join(timeout=120)
  assert results == {0: "ok", 1: "ok"}, results


def test_two_rank_training_step_allreduce():
  _run(_train_worker)


def test_two_rank_sharded_codec_strings():
  _run(_codec_worker)
