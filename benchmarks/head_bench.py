#!/usr/bin/env python
"""Times the heads kernel (experiment.py:199-203 over [T*B, 256], :237) and the
learner block from the core output on.

    python benchmarks/head_bench.py [--out gpurun_out/head.jsonl]

One JSON line per (shape, implementation):
* `head tcgen05`  vtr_policy_head_f32 (3xTF32 tcgen05 kernel of this library);
  HBM-bound: 4K bytes read + 4(A+1) written per row - `GBps` is on those bytes.
* `head cuBLAS fp32`  torch.addmm on the concatenated [K, A+1] weights with
  TF32 off (the library contraction the reference's snt.Linear amounts to).
* `head cuBLAS tf32`  the same with allow_tf32 (1e-3 accurate: shown for
  speed only, it does not meet the path's 1e-5).
* `block: ...`  heads + fused V-trace/loss forward+backward pass, with either
  head implementation in front of the same fused pass.
Timing: CUDA events around back-to-back calls after warm-up; the core output
of the large shapes (3.4 GB) is far beyond L2.
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402
from scalable_agent_b200 import _lib, heads  # noqa: E402


def timed(fn, iters, warm=5):
  for _ in range(warm):
    fn()
  torch.cuda.synchronize()
  s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  s.record()
  for _ in range(iters):
    fn()
  e.record()
  torch.cuda.synchronize()
  return s.elapsed_time(e) * 1e3 / iters


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--out', default=os.path.join(ROOT, 'gpurun_out', 'head.jsonl'))
  args = ap.parse_args()
  device = torch.device('cuda', 0)
  torch.cuda.set_device(0)
  lib = _lib.load()
  peak, _ = bench.measured_peak_gbs()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  os.makedirs(os.path.dirname(args.out), exist_ok=True)
  K = 256
  lines = []
  for name, T, B, A in [('config2', 100, 32, 9), ('config3', 100, 2048, 9), ('sweep_B4096', 100, 4096, 18),
                        ('target_B32768', 100, 32768, 18)]:
    rows = T * B
    g = torch.Generator(device=device).manual_seed(7)
    core = torch.randn(rows, K, generator=g, device=device)
    wp = torch.randn(K, A, generator=g, device=device) / 16
    bp = torch.randn(A, generator=g, device=device)
    wv = torch.randn(K, generator=g, device=device) / 16
    bv = torch.randn(1, generator=g, device=device)
    packed = heads.pack_head_weights(wp, bp, wv, bv)
    logits = torch.empty(rows, A, device=device)
    baseline = torch.empty(rows, device=device)
    w_cat = torch.cat([wp, wv[:, None]], 1).contiguous()
    b_cat = torch.cat([bp, bv]).contiguous()
    out_cat = torch.empty(rows, A + 1, device=device)
    nbytes = rows * (4 * K + 4 * (A + 1)) + 4 * K * (A + 1)

    def ours():
      rc = lib.vtr_policy_head_f32(core.data_ptr(), packed.data_ptr(), rows, K, A, logits.data_ptr(),
                                   baseline.data_ptr(), stream)
      assert rc == 0, _lib.last_error()

    def cublas():
      torch.addmm(b_cat, core, w_cat, out=out_cat)

    iters = 200 if rows < 1e6 else 30
    torch.backends.cuda.matmul.allow_tf32 = False
    res = [('head tcgen05 3xTF32', timed(ours, iters)), ('head cuBLAS fp32', timed(cublas, iters))]
    torch.backends.cuda.matmul.allow_tf32 = True
    res.append(('head cuBLAS tf32 (1e-3 accurate)', timed(cublas, iters)))
    torch.backends.cuda.matmul.allow_tf32 = False
    # accuracy of each against float64 on a sample of rows
    sample = slice(0, min(rows, 4096))
    ref = (core[sample].double() @ w_cat.double() + b_cat.double())
    ours()
    err_ours = float(((torch.cat([logits[sample], baseline[sample, None]], 1).double() - ref).abs() /
                      ref.abs().clamp(min=1)).max())
    cublas()
    err_cublas = float(((out_cat[sample].double() - ref).abs() / ref.abs().clamp(min=1)).max())

    # the block from the core output on: heads + fused pass (forward + backward of the loss)
    inp = bench.synthetic_on_device(T, B, A, 1, device)
    call = bench.FusedCall(inp, T, B, A, device)
    call.p.target_policy_logits = logits.data_ptr()
    call.p.values = baseline.data_ptr()

    def block_ours():
      ours()
      call()

    def block_cublas():
      cublas()                      # logits and baseline interleaved [rows, A+1]: split them as the fused pass wants
      logits.copy_(out_cat[:, :A])
      baseline.copy_(out_cat[:, A])
      call()

    res += [('block: tcgen05 heads + fused pass', timed(block_ours, iters)),
            ('block: cuBLAS fp32 heads + split + fused pass', timed(block_cublas, iters))]
    for label, us in res:
      rec = {'shape': name, 'T': T, 'B': B, 'A': A, 'K': K, 'impl': label, 'us': us}
      if label.startswith('head'):
        rec.update(algorithmic_bytes=nbytes, GBps=nbytes / us / 1e3, frac_peak=nbytes / us / 1e3 / peak)
      if label.startswith('head tcgen05'):
        rec['max_scaled_err_vs_f64'] = err_ours
      if label == 'head cuBLAS fp32':
        rec['max_scaled_err_vs_f64'] = err_cublas
      lines.append(rec)
      print('%-14s %-46s %9.1f us %s' % (name, label, us,
                                         '%7.0f GB/s (%.2f of peak)' % (rec['GBps'], rec['frac_peak'])
                                         if 'GBps' in rec else ''))
    print('%-14s max scaled error vs float64: tcgen05 %.2e, cuBLAS fp32 %.2e' % (name, err_ours, err_cublas))
    del core, inp, call, out_cat, logits, baseline
    torch.cuda.empty_cache()
  with open(args.out, 'w') as f:
    for rec in lines:
      f.write(json.dumps(rec) + '\n')


if __name__ == '__main__':
  main()
