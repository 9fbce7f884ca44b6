"""The NVLink peer-memory loss exchange (vtr_exchange_*), on 2 GPUs.

Needs two CUDA devices; skipped on a one-GPU box (the driver's `-m gpu` run).
Run with: gpurun --gpus 2 -- python -m pytest tests/test_exchange_gpu.py -m gpu
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
  with socket.socket() as s:
    s.bind(('127.0.0.1', 0))
    return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  torch.cuda.set_device(rank)
  device = torch.device('cuda', rank)
  dist.init_process_group('nccl', rank=rank, world_size=world, device_id=device)
  try:
    from scalable_agent_b200.distributed import LossExchange
    ex = LossExchange(device)
    results = []
    for step in range(1, 6):
      mine = torch.tensor([rank + 0.25 * step, -3.0 * rank, 1e-3 * step, 7.0],
                          device=device)
      results.append(ex.all_sum(mine).cpu())
    torch.save(torch.stack(results), os.path.join(out_dir, 'r%d.pt' % rank))
  finally:
    dist.destroy_process_group()


def test_loss_exchange_two_ranks(tmp_path):
  if torch.cuda.device_count() < 2:
    pytest.skip('needs 2 GPUs')
  world = 2
  mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world,
           join=True)
  r0 = torch.load(tmp_path / 'r0.pt')
  r1 = torch.load(tmp_path / 'r1.pt')
  assert torch.equal(r0, r1), 'every rank must hold the same bits'
  for i, step in enumerate(range(1, 6)):
    want = torch.tensor([(0 + 0.25 * step) + (1 + 0.25 * step), -3.0,
                         2e-3 * step, 14.0])
    torch.testing.assert_close(r0[i], want, rtol=1e-6, atol=1e-6)
