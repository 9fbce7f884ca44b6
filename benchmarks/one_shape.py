#!/usr/bin/env python
"""Runs the fused pass N times on one shape / kernel path: the short command
line that ncu captures are taken on (see profiles/README.md).

    python benchmarks/one_shape.py T B A [path] [calls]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from scalable_agent_b200 import _lib  # noqa: E402


def main():
  T, B, A = (int(x) for x in sys.argv[1:4])
  path = sys.argv[4] if len(sys.argv) > 4 else 'auto'
  calls = int(sys.argv[5]) if len(sys.argv) > 5 else 30
  device = torch.device('cuda', 0)
  torch.cuda.set_device(0)
  call = bench.FusedCall(bench.synthetic_on_device(T, B, A, 1, device), T, B, A, device)
  call.p.path = _lib.PATHS[path]
  print(call.describe())
  for _ in range(calls):
    call()
  torch.cuda.synchronize()
  print('losses', call.out['losses'].tolist())


if __name__ == '__main__':
  main()
