#!/usr/bin/env python
"""Runs the heads kernel a few times on one shape: the short command line that
ncu captures are taken on.   python benchmarks/one_head.py [rows] [A] [calls]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from scalable_agent_b200 import _lib, heads  # noqa: E402


def main():
  rows = int(sys.argv[1]) if len(sys.argv) > 1 else 3276800
  A = int(sys.argv[2]) if len(sys.argv) > 2 else 18
  calls = int(sys.argv[3]) if len(sys.argv) > 3 else 10
  K = 256
  dev = torch.device('cuda', 0)
  g = torch.Generator(device=dev).manual_seed(1)
  core = torch.randn(rows, K, generator=g, device=dev)
  packed = heads.pack_head_weights(torch.randn(K, A, generator=g, device=dev) / 16, torch.zeros(A, device=dev),
                                   torch.randn(K, generator=g, device=dev) / 16, torch.zeros(1, device=dev))
  logits = torch.empty(rows, A, device=dev)
  baseline = torch.empty(rows, device=dev)
  lib = _lib.load()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  for i in range(calls):
    if i == calls // 2:
      start.record()
    rc = lib.vtr_policy_head_f32(core.data_ptr(), packed.data_ptr(), rows, K, A, logits.data_ptr(),
                                 baseline.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
  end.record()
  torch.cuda.synchronize()
  print('ok %.1f us per call' % (start.elapsed_time(end) * 1e3 / (calls - calls // 2)),
        float(logits[0, 0]), float(baseline[-1]))


if __name__ == '__main__':
  main()
