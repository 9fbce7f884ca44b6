#!/usr/bin/env python
"""Times the fused pass at every BASELINE.json configuration (1 GPU).

    python benchmarks/sweep.py [--out gpurun_out/sweep.jsonl] [--paths auto,column,scan]

For each shape: `cold` = median CUDA-event time of single calls with the L2
flushed (a 512 MB buffer is overwritten) before each one; `warm` = mean over
back-to-back calls on the same inputs (inputs of the small shapes then live in
L2, as they would right after the network's heads produced them).  One JSON
line per (shape, path).  This is a diagnostic; `bench.py` is the contract.
"""
import argparse
import ctypes
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402

SHAPES = [
    # (name, T, B, A)
    ('config1_plumbing', 5, 2, 3),
    ('config2_dmlab_single', 100, 32, 9),
    ('config3_dmlab30_large', 100, 2048, 9),
    ('config4_short_unroll', 20, 16384, 18),
    ('sweep_B4096', 100, 4096, 18),
    ('sweep_B8192', 100, 8192, 18),
    ('sweep_B16384', 100, 16384, 18),
    ('target_B32768', 100, 32768, 18),
    ('sweep_B65536', 100, 65536, 18),
    ('sweep_B131072', 100, 131072, 18),
    ('sweep_B262144', 100, 262144, 18),
]


def main():
  import torch
  from scalable_agent_b200 import _lib
  ap = argparse.ArgumentParser()
  ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'sweep.jsonl'))
  ap.add_argument('--paths', default='auto')
  ap.add_argument('--only', default='')
  ap.add_argument('--extra', default='', help='extra shapes: TxBxA,TxBxA')
  ap.add_argument('--long', action='store_true', help='20x more warm iterations')
  args = ap.parse_args()
  device = torch.device('cuda', 0)
  torch.cuda.set_device(0)
  peak, _ = bench.measured_peak_gbs()
  flush = torch.empty(512 * 1024 * 1024 // 4, dtype=torch.float32, device=device)
  os.makedirs(os.path.dirname(args.out), exist_ok=True)
  shapes = list(SHAPES)
  for spec in filter(None, args.extra.split(',')):
    T, B, A = (int(x) for x in spec.split('x'))
    shapes.append(('extra_%s' % spec, T, B, A))
  with open(args.out, 'w') as f:
    for name, T, B, A in shapes:
      if args.only and not any(tag in name for tag in args.only.split(',')):
        continue
      inp = bench.synthetic_on_device(T, B, A, 1, device)
      for path in args.paths.split(','):
        call = bench.FusedCall(inp, T, B, A, device)
        call.p.path = _lib.PATHS[path]
        try:
          call.describe()
          for _ in range(5):
            call()
          torch.cuda.synchronize()
        except (ValueError, RuntimeError) as e:
          print('%s path=%s: %s' % (name, path, e))
          continue
        cold = []
        for _ in range(15):
          flush.fill_(1.0)
          s = torch.cuda.Event(enable_timing=True)
          e = torch.cuda.Event(enable_timing=True)
          s.record()
          call()
          e.record()
          torch.cuda.synchronize()
          cold.append(s.elapsed_time(e) * 1e3)
        iters = 200 if T * B < 4e6 else 50
        if args.long:
          iters *= 20
        sampler = bench.ClockSampler(bench.physical_gpu_index(0))
        sampler.start()
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
          call()
        e.record()
        torch.cuda.synchronize()
        sampler.stop()
        clocks = sampler.summary()
        warm_us = s.elapsed_time(e) * 1e3 / iters
        # The same calls replayed from a CUDA graph: no Python, ctypes or
        # launch API between the kernels, so this is what the GPU itself needs
        # per call (for the small shapes `warm` is the host's launch rate).
        graph_us = None
        try:
          per_graph = 20
          side = torch.cuda.Stream()
          side.wait_stream(torch.cuda.current_stream())
          graph = torch.cuda.CUDAGraph()
          with torch.cuda.graph(graph, stream=side):
            cap = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            for _ in range(per_graph):
              call(cap)
          for _ in range(3):
            graph.replay()
          torch.cuda.synchronize()
          reps = max(2, iters // per_graph)
          s.record()
          for _ in range(reps):
            graph.replay()
          e.record()
          torch.cuda.synchronize()
          graph_us = s.elapsed_time(e) * 1e3 / (reps * per_graph)
          del graph
        except Exception as err:  # pylint: disable=broad-except
          print('%s path=%s: no graph timing: %s' % (name, path, str(err).splitlines()[0][:120]))
        alg = bench.algorithmic_bytes(T, B, A)
        cold_us = statistics.median(cold)
        rec = {
            'shape': name, 'T': T, 'B': B, 'A': A, 'path': path,
            'cold_us': cold_us, 'warm_us': warm_us, 'graph_us': graph_us,
            'plan': call.plan,
            'timesteps_per_s_warm': T * B / (warm_us * 1e-6),
            'GBps_cold': alg / cold_us / 1e3, 'GBps_warm': alg / warm_us / 1e3,
            'frac_peak_warm': alg / warm_us / 1e3 / peak,
            'roofline_us': alg / peak / 1e3, 'clocks': clocks,
        }
        f.write(json.dumps(rec) + '\n')
        f.flush()
        print('%-24s path=%-6s cold %8.1f us  warm %8.1f us  graph %s us  %7.0f GB/s (%.2f of peak) '
              'roofline %.1f us  sm %s MHz %s' % (name, path, cold_us, warm_us,
                                    '%8.1f' % graph_us if graph_us else '     n/a', rec['GBps_warm'],
                                    rec['frac_peak_warm'], rec['roofline_us'],
                                    clocks['sm_mhz'], ','.join(clocks['reasons'])))
      del inp
      torch.cuda.empty_cache()


if __name__ == '__main__':
  main()
