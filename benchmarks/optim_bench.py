#!/usr/bin/env python
"""Times the optimiser step that follows the path (experiment.py:410-415).

    python benchmarks/optim_bench.py                      # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port P benchmarks/optim_bench.py

One JSON line per (parameter count, implementation):

* `rmsprop` (1 GPU): vtr_rmsprop_f32, HBM-bound - 16 B read + 12 B written per
  parameter with the momentum slot, 12 + 8 without (the reference's momentum
  is 0, experiment.py:94).  `gbs` is on those algorithmic bytes.
* `fused` (N GPUs): vtr_rmsprop_allreduce_f32, one launch per rank doing
  reduce-scatter + update + all-gather over NVLink peer memory.
* `nccl` (N GPUs): the same step from library parts, timed beside it as the
  baseline: ncclAllReduce(SUM) on the gradient, then vtr_rmsprop_f32 on the
  full parameter vector on every rank.

Sizes: 1.6 M parameters (the order of the reference's shallow agent) up to
64 M.  Timing: CUDA events around K back-to-back steps after W warm-ups, max
over ranks.  A diagnostic; `bench.py` is the contract.
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LR, DECAY, MOMENTUM, EPSILON = 0.00048, 0.99, 0.0, 0.1   # experiment.py:92-95


def timed(fn, steps, warmup, barrier):
  import torch
  for _ in range(warmup):
    fn()
  barrier()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(steps):
    fn()
  b.record()
  barrier()
  return a.elapsed_time(b) / steps * 1e3     # us per step


def main():
  import torch
  import torch.distributed as dist
  from scalable_agent_b200 import _lib, optim

  ap = argparse.ArgumentParser()
  ap.add_argument('--sizes', default='1600000,16000000,64000000')
  ap.add_argument('--steps', type=int, default=100)
  ap.add_argument('--warmup', type=int, default=10)
  ap.add_argument('--out', default='')
  args = ap.parse_args()

  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local_rank = int(os.environ.get('LOCAL_RANK', '0'))
  torch.cuda.set_device(local_rank)
  device = torch.device('cuda', local_rank)
  if world > 1:
    dist.init_process_group('nccl', device_id=device)
  lib = _lib.load()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  def max_over_ranks(us):
    if world == 1:
      return us
    t = torch.tensor([us], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

  lines = []
  for n in [int(s) for s in args.sizes.split(',')]:
    if world == 1:
      for with_mom in (False, True):
        var = torch.randn(n, device=device)
        ms = torch.ones(n, device=device)
        mom = torch.zeros(n, device=device) if with_mom else None
        grad = torch.randn(n, device=device)

        def step():
          rc = lib.vtr_rmsprop_f32(var.data_ptr(), ms.data_ptr(),
                                   mom.data_ptr() if mom is not None else None,
                                   grad.data_ptr(), n, LR, DECAY,
                                   0.9 if with_mom else 0.0, EPSILON, stream)
          assert rc == 0, _lib.last_error()
        us = timed(step, args.steps, args.warmup, barrier)
        nbytes = n * (28 if with_mom else 20)
        lines.append({'impl': 'rmsprop', 'n_gpus': 1, 'params': n, 'momentum_slot': with_mom,
                      'us_per_step': us, 'algorithmic_bytes': nbytes,
                      'gbs': nbytes / us * 1e-3})
    else:
      params = [torch.nn.Parameter(torch.randn(n, device=device))]
      opt = optim.RMSPropOptimizer(params, LR, DECAY, MOMENTUM, EPSILON, group=dist.group.WORLD)
      # optim_check: a few steps of the fused kernel against the same steps
      # from library parts (ncclAllReduce of the gradients in rank order is
      # not guaranteed, so the reference is an all-gather + rank-ordered sum,
      # then vtr_rmsprop_f32); parameters must be the same bits on every rank
      # and equal to the reference.
      ref_var = params[0].detach().clone()
      ref_ms = torch.ones(n, device=device)
      gathered = torch.empty(world * n, device=device) if n <= 16000000 else None
      check = {'ok': None, 'reason': 'skipped above 16 M parameters'}
      if gathered is not None:
        gen = torch.Generator(device=device).manual_seed(1234 + rank)
        ok = True
        for _ in range(3):
          params[0].grad.copy_(torch.randn(n, generator=gen, device=device))
          dist.all_gather_into_tensor(gathered, params[0].grad)
          total = gathered[:n].clone()
          for r in range(1, world):
            total += gathered[r * n:(r + 1) * n]
          opt.step()
          rc = lib.vtr_rmsprop_f32(ref_var.data_ptr(), ref_ms.data_ptr(), None, total.data_ptr(),
                                   n, LR, DECAY, 0.0, EPSILON, stream)
          assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        opt.check_peers()
        same_as_ref = torch.equal(params[0].detach(), ref_var)
        mine = params[0].detach().clone()
        dist.broadcast(mine, src=0)
        same_on_all = torch.equal(mine, params[0].detach())
        flag = torch.tensor([1 if (same_as_ref and same_on_all) else 0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        check = {'ok': bool(int(flag.item())), 'steps': 3,
                 'max_abs_diff': float((params[0].detach() - ref_var).abs().max().item()),
                 'against': 'all_gather + rank-ordered sum + vtr_rmsprop_f32, bitwise; '
                            'identical parameters on all ranks'}
        del gathered
      params[0].grad.normal_()
      us = max_over_ranks(timed(opt.step, args.steps, args.warmup, barrier))
      opt.check_peers()
      # into a rank: what it loads from its N-1 peers (their part of the sum for
      # its slice) plus what the N-1 peers store into its parameters (their
      # slices); out of it: the mirror image
      per_direction = 2 * (opt.hi - opt.lo) * 4 * (world - 1)
      lines.append({'impl': 'fused', 'n_gpus': world, 'params': n, 'us_per_step': us,
                    'optim_check': check,
                    'nvlink_bytes_per_direction_per_rank': per_direction,
                    'nvlink_gbs_per_direction': per_direction / us * 1e-3})
      del opt, params
      var = torch.randn(n, device=device)
      ms = torch.ones(n, device=device)
      grad = torch.randn(n, device=device)

      def nccl_step():
        dist.all_reduce(grad)
        rc = lib.vtr_rmsprop_f32(var.data_ptr(), ms.data_ptr(), None, grad.data_ptr(), n,
                                 LR, DECAY, 0.0, EPSILON, stream)
        assert rc == 0, _lib.last_error()
      us = max_over_ranks(timed(nccl_step, args.steps, args.warmup, barrier))
      lines.append({'impl': 'nccl', 'n_gpus': world, 'params': n, 'us_per_step': us})
  if rank == 0:
    text = '\n'.join(json.dumps(l) for l in lines)
    print(text, flush=True)
    if args.out:
      with open(args.out, 'w') as f:
        f.write(text + '\n')
  if world > 1:
    dist.barrier()
    dist.destroy_process_group()
  if any(l.get('optim_check', {}).get('ok') is False for l in lines):
    raise SystemExit('optim_bench.py: the fused all-reduce + RMSProp step does not match '
                     'the reference')


if __name__ == '__main__':
  main()
