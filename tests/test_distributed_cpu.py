"""Host-side logic of the B-sharded path, on CPU with gloo and world_size 2.

What is checked: contiguous balanced shards; the rank-ordered deterministic
sum of the loss scalars gives identical bits on every rank; and the property
the sharding relies on -- per-column V-trace results and the summed losses of
the shards equal those of the unsharded batch (here through the oracle, since
the CUDA kernels cannot run on CPU).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import vtrace_oracle as oracle
from scalable_agent_b200 import distributed as vdist

KEYS = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
        'discounts', 'rewards', 'values', 'bootstrap_value')


def test_shard_bounds_cover_and_balance():
  for B in (1, 7, 32, 33, 32768):
    for world in (1, 2, 3, 8):
      bounds = [vdist.shard_bounds(B, world, r) for r in range(world)]
      assert bounds[0][0] == 0 and bounds[-1][1] == B
      for (lo, hi), (lo2, _) in zip(bounds, bounds[1:]):
        assert hi == lo2
      sizes = [hi - lo for lo, hi in bounds]
      assert max(sizes) - min(sizes) <= 1
  with pytest.raises(ValueError):
    vdist.shard_bounds(8, 2, 2)


def test_all_reduce_losses_without_process_group():
  x = torch.tensor([1.0, 2.0, 3.0, 4.0])
  assert torch.equal(vdist.all_reduce_losses(x), x)


def _free_port():
  with socket.socket() as s:
    s.bind(('127.0.0.1', 0))
    return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  dist.init_process_group('gloo', rank=rank, world_size=world)
  try:
    T, B, A = 12, 10, 5
    inp = oracle.synthetic_inputs(T, B, A, seed=3)
    lo, hi = vdist.shard_bounds(B, world, rank)
    shard = {}
    for k in KEYS:
      x = torch.from_numpy(inp[k])
      shard[k] = vdist.shard_time_major(x, world, rank,
                                        batch_dim=0 if x.dim() == 1 else 1)
      assert shard[k].is_contiguous()
    ref = oracle.fused_vtrace_loss(*[shard[k].numpy() for k in KEYS])
    local = torch.tensor([ref.total_loss, ref.pg_loss, ref.baseline_loss,
                          ref.entropy_loss], dtype=torch.float32)
    total = vdist.all_reduce_losses(local)
    np.save(os.path.join(out_dir, 'losses_%d.npy' % rank), total.numpy())
    np.save(os.path.join(out_dir, 'vs_%d.npy' % rank), ref.vs)
    np.save(os.path.join(out_dir, 'bounds_%d.npy' % rank), np.array([lo, hi]))
  finally:
    dist.destroy_process_group()


def test_two_rank_sharding_matches_unsharded(tmp_path):
  world = 2
  mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world,
           join=True)
  T, B, A = 12, 10, 5
  inp = oracle.synthetic_inputs(T, B, A, seed=3)
  full = oracle.fused_vtrace_loss(*[inp[k] for k in KEYS])
  l0 = np.load(tmp_path / 'losses_0.npy')
  l1 = np.load(tmp_path / 'losses_1.npy')
  assert np.array_equal(l0, l1), 'ranks must hold identical bits'
  want = np.array([full.total_loss, full.pg_loss, full.baseline_loss,
                   full.entropy_loss])
  np.testing.assert_allclose(l0, want, rtol=1e-6)
  for r in range(world):
    lo, hi = np.load(tmp_path / ('bounds_%d.npy' % r))
    np.testing.assert_allclose(np.load(tmp_path / ('vs_%d.npy' % r)),
                               full.vs[:, lo:hi], rtol=0, atol=1e-12)
