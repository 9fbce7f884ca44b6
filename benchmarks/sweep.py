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


def surface(args, device, peak):
  """The drop-in surface a caller that does NOT use the fused entry point
  sees: each stand-alone function of the reference (vtrace.py:45,164;
  experiment.py:324-343) on its own, and the composed reference-style call
  (from_logits + the three helpers + autograd backward), against each
  kernel's own algorithmic bytes.  One JSON line per (shape, op)."""
  import torch
  from scalable_agent_b200 import _ops, experiment, vtrace

  def timed(fn, iters):
    for _ in range(5):
      fn()
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True)
    e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
      fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e3 / iters

  shapes = [('config2_dmlab_single', 100, 32, 9), ('config3_dmlab30_large', 100, 2048, 9),
            ('sweep_B4096', 100, 4096, 18), ('target_B32768', 100, 32768, 18)]
  with open(args.out, 'w') as f:
    for name, T, B, A in shapes:
      if args.only and not any(tag in name for tag in args.only.split(',')):
        continue
      inp = bench.synthetic_on_device(T, B, A, 1, device)
      rows = T * B
      logits, actions = inp['target_policy_logits'], inp['actions']
      adv = torch.randn(T, B, device=device)
      one = torch.ones(1, device=device)
      log_rhos = 0.3 * torch.randn(T, B, device=device)
      C = 42
      rew_c = torch.randn(T, B, C, device=device)
      val_c = torch.randn(T, B, C, device=device)
      boot_c = torch.randn(B, C, device=device)
      ops = [
          # (label, bytes per call, callable)
          ('log_probs_from_logits_and_actions', rows * (4 * A + 8),
           lambda: _ops.call(_ops.log_probs, logits, actions)),
          ('compute_policy_gradient_loss fwd', rows * (4 * A + 8),
           lambda: _ops.call(_ops.pg_loss, logits, actions, adv)),
          ('compute_policy_gradient_loss bwd', rows * (8 * A + 8),
           lambda: _ops.call(_ops.pg_loss_bwd, logits, actions, adv, one)),
          ('compute_entropy_loss fwd', rows * 4 * A,
           lambda: _ops.call(_ops.entropy_loss, logits)),
          ('compute_entropy_loss bwd', rows * 8 * A,
           lambda: _ops.call(_ops.entropy_loss_bwd, logits, one)),
          ('compute_baseline_loss fwd', rows * 4,
           lambda: _ops.call(_ops.baseline_loss, adv)),
          ('compute_baseline_loss bwd', rows * 8,
           lambda: _ops.call(_ops.baseline_loss_bwd, adv, one)),
          ('from_importance_weights [T,B]', rows * 24 + 4 * B,
           lambda: vtrace.from_importance_weights(log_rhos, inp['discounts'], inp['rewards'],
                                                  inp['values'], inp['bootstrap_value'])),
          ('from_importance_weights [T,B], op only', rows * 24 + 4 * B,
           lambda: _ops.call(_ops.from_importance_weights, log_rhos, inp['discounts'], inp['rewards'],
                             inp['values'].reshape(T, B, 1), inp['bootstrap_value'], 1, 0, 1.0, 1.0)),
          ('from_importance_weights [T,B,1]x[T,B,42]', rows * (8 + 16 * C) + 4 * B * C,
           lambda: vtrace.from_importance_weights(log_rhos[..., None], inp['discounts'][..., None],
                                                  rew_c, val_c, boot_c)),
          ('from_logits (forward only, 5 fields)', rows * (8 * A + 36) + 4 * B,
           lambda: vtrace.from_logits(*[inp[k] for k in bench.FUSED_KEYS])),
      ]
      tl = inp['target_policy_logits'].detach().requires_grad_(True)
      vl = inp['values'].detach().requires_grad_(True)

      def composed():
        # the reference's learner block, function by function (experiment.py:389-407, :415)
        tl.grad = None
        vl.grad = None
        vt = vtrace.from_logits(inp['behaviour_policy_logits'], tl, inp['actions'], inp['discounts'],
                                inp['rewards'], vl, inp['bootstrap_value'])
        total = experiment.compute_policy_gradient_loss(tl, inp['actions'], vt.pg_advantages)
        total = total + 0.5 * experiment.compute_baseline_loss(vt.vs - vl)
        total = total + 0.00025 * experiment.compute_entropy_loss(tl)
        total.backward()

      def fused():
        tl.grad = None
        vl.grad = None
        experiment.compute_vtrace_loss(inp['behaviour_policy_logits'], tl, inp['actions'],
                                       inp['discounts'], inp['rewards'], vl,
                                       inp['bootstrap_value']).total_loss.backward()

      fused_bytes = bench.algorithmic_bytes(T, B, A)
      ops += [('composed: from_logits + 3 helpers + backward', fused_bytes, composed),
              ('fused: compute_vtrace_loss + backward', fused_bytes, fused)]
      iters = 300 if rows < 1e6 else 60
      for label, nbytes, fn in ops:
        us = timed(fn, iters)
        rec = {'shape': name, 'T': T, 'B': B, 'A': A, 'op': label, 'us': us,
               'algorithmic_bytes': nbytes, 'GBps': nbytes / us / 1e3,
               'frac_peak': nbytes / us / 1e3 / peak, 'timesteps_per_s': rows / (us * 1e-6),
               'note': 'bytes of the fused pass' if label.startswith(('composed', 'fused')) else ''}
        f.write(json.dumps(rec) + '\n')
        f.flush()
        print('%-22s %-46s %9.1f us  %7.0f GB/s (%.2f of peak)' % (name, label, us, rec['GBps'],
                                                                   rec['frac_peak']))


def main():
  import torch
  from scalable_agent_b200 import _lib
  ap = argparse.ArgumentParser()
  ap.add_argument('--surface', action='store_true',
                  help='time the stand-alone functions and the composed call instead')
  ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'sweep.jsonl'))
  ap.add_argument('--paths', default='auto')
  ap.add_argument('--only', default='')
  ap.add_argument('--extra', default='', help='extra shapes: TxBxA,TxBxA')
  ap.add_argument('--long', action='store_true', help='20x more warm iterations')
  args = ap.parse_args()
  device = torch.device('cuda', 0)
  torch.cuda.set_device(0)
  peak, _ = bench.measured_peak_gbs()
  os.makedirs(os.path.dirname(args.out), exist_ok=True)
  if args.surface:
    surface(args, device, peak)
    return
  flush = torch.empty(512 * 1024 * 1024 // 4, dtype=torch.float32, device=device)
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
