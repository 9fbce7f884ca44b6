#!/usr/bin/env python
"""GPU timeline of a few `compute_vtrace_loss(...)` + `backward()` steps
(torch.profiler / CUPTI): every kernel and memset with its start time and
duration, so that idle gaps between them show.  A diagnostic."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402
from scalable_agent_b200 import experiment  # noqa: E402


def main():
  T, B, A = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (100, 32768, 18)
  device = torch.device('cuda', 0)
  inp = bench.synthetic_on_device(T, B, A, 1, device)
  inp['target_policy_logits'].requires_grad_(True)
  inp['values'].requires_grad_(True)

  def step():
    inp['target_policy_logits'].grad = None
    inp['values'].grad = None
    experiment.compute_vtrace_loss(baseline_cost=0.5, entropy_cost=0.00025,
                                   **inp).total_loss.backward()

  for _ in range(30):
    step()
  torch.cuda.synchronize()
  with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(6):
      step()
    torch.cuda.synchronize()
  events = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
  events.sort(key=lambda e: e.time_range.start)
  t0 = events[0].time_range.start
  prev_end = t0
  for e in events:
    start, end = e.time_range.start, e.time_range.end
    print('%9.1f us  +%6.1f gap  %8.1f us  %s' % (start - t0, start - prev_end, end - start, e.name[:90]))
    prev_end = max(prev_end, end)


if __name__ == '__main__':
  main()
