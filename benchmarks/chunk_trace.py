#!/usr/bin/env python
"""Where a K5 CTA spends its time: per-phase time stamps (debug build only).

    make -C scalable_agent_b200/csrc trace      # builds lib/libvtrace_b200_trace.so
    python benchmarks/chunk_trace.py [T B A]

Prints, over all CTAs of one warm launch, the median / max time (us, from the
earliest CTA start) at which each phase was reached.
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ['VTRACE_B200_LIB'] = os.path.join(
    ROOT, 'scalable_agent_b200', 'lib', 'libvtrace_b200_trace.so')

import numpy as np   # noqa: E402
import torch         # noqa: E402

import bench         # noqa: E402
from scalable_agent_b200 import _lib   # noqa: E402

NAMES = ['start', 'loads issued', 'rows landed', 'phase A done', 'cluster sync done',
         'rescan done', 'phase C done', 'end']


def main():
  T, B, A = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (100, 2048, 9)
  device = torch.device('cuda', 0)
  inp = bench.synthetic_on_device(T, B, A, 1, device)
  trace = torch.zeros(8 * 65536, dtype=torch.int64, device=device)
  os.environ['VTR_CHUNK_TRACE_PTR'] = '%x' % trace.data_ptr()
  call = bench.FusedCall(inp, T, B, A, device)
  call.p.path = _lib.PATHS['chunk']
  print(call.plan)
  for _ in range(5):
    call()
  torch.cuda.synchronize()
  trace.zero_()
  call()
  torch.cuda.synchronize()
  t = trace.cpu().numpy().reshape(-1, 8)
  t = t[t[:, 0] > 0].astype(np.float64)
  t0 = t[:, 0].min()
  print('%d CTAs' % len(t))
  for k, name in enumerate(NAMES):
    rel = (t[:, k] - t0) / 1e3
    own = (t[:, k] - t[:, 0]) / 1e3
    print('%-18s since kernel start: median %6.2f max %6.2f us | since CTA start: median %6.2f max %6.2f us'
          % (name, np.median(rel), rel.max(), np.median(own), own.max()))


if __name__ == '__main__':
  main()
