#!/usr/bin/env python
"""Host cost of the Python drop-in: where the microseconds of
`experiment.compute_vtrace_loss(...)` + `backward()` go (cProfile) and the
host time of the forward and the backward half per call.

    python benchmarks/api_profile.py [T B A]

A diagnostic for DESIGN.md section 1; `bench.py`'s `api_device` leg is the
number that counts.
"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from scalable_agent_b200 import experiment  # noqa: E402


def main():
  T, B, A = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (100, 32, 9)
  device = torch.device('cuda', 0)
  inp = bench.synthetic_on_device(T, B, A, 1, device)
  inp['target_policy_logits'].requires_grad_(True)
  inp['values'].requires_grad_(True)

  def fwd():
    return experiment.compute_vtrace_loss(baseline_cost=0.5, entropy_cost=0.00025, **inp)

  def step():
    inp['target_policy_logits'].grad = None
    inp['values'].grad = None
    fwd().total_loss.backward()

  for _ in range(50):
    step()
  torch.cuda.synchronize()
  n = 500
  t_fwd = t_bwd = 0.0
  for _ in range(n):
    inp['target_policy_logits'].grad = None
    inp['values'].grad = None
    t0 = time.perf_counter()
    out = fwd()
    t1 = time.perf_counter()
    out.total_loss.backward()
    t2 = time.perf_counter()
    t_fwd += t1 - t0
    t_bwd += t2 - t1
    del out
  torch.cuda.synchronize()
  print('T=%d B=%d A=%d: host time per call: forward %.1f us, backward %.1f us' % (
      T, B, A, t_fwd / n * 1e6, t_bwd / n * 1e6))
  # the autograd engine hands CUDA nodes to a device thread; run inline instead
  with torch.autograd.set_multithreading_enabled(False):
    t_bwd = 0.0
    for _ in range(n):
      inp['target_policy_logits'].grad = None
      inp['values'].grad = None
      out = fwd()
      t1 = time.perf_counter()
      out.total_loss.backward()
      t_bwd += time.perf_counter() - t1
      del out
    torch.cuda.synchronize()
    print('backward with torch.autograd.set_multithreading_enabled(False): %.1f us' % (t_bwd / n * 1e6))
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
      step()
    b.record()
    torch.cuda.synchronize()
    print('device time per forward+backward, engine inline: %.1f us' % (a.elapsed_time(b) * 1e3 / n))
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(n):
    step()
  b.record()
  torch.cuda.synchronize()
  print('device time per forward+backward: %.1f us' % (a.elapsed_time(b) * 1e3 / n))
  prof = cProfile.Profile()
  prof.enable()
  for _ in range(n):
    step()
  prof.disable()
  torch.cuda.synchronize()
  stats = pstats.Stats(prof)
  stats.sort_stats('tottime').print_stats(16)


if __name__ == '__main__':
  main()
