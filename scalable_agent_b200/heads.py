"""The policy and baseline heads that produce the path's inputs.

`/root/reference/experiment.py:199-203`:

    policy_logits = snt.Linear(self._num_actions, name='policy_logits')(core_output)
    baseline = tf.squeeze(snt.Linear(1, name='baseline')(core_output), axis=-1)

applied to the `[T*B, 256]` core output by `snt.BatchApply(self._head)` (`:237`).
This is the one dense contraction next to the V-trace path; here it is one
`tcgen05` tensor-core kernel (`csrc/vtrace_head.cu`) that computes both heads
in a single pass over the core output with fp32-accurate (3xTF32) products, and
writes `policy_logits` / `baseline` in the layout the fused pass reads.

The forward is this library's kernel.  The backward (two more contractions,
`d_core = d_out W^T` and `dW = core^T d_out`) is left to torch's matmul.
"""
import ctypes

import torch

from scalable_agent_b200 import _lib
from scalable_agent_b200 import _ops


class _HeadFn(torch.autograd.Function):

  @staticmethod
  def forward(ctx, core, w_policy, b_policy, w_baseline, b_baseline):
    logits, baseline = policy_head_forward(core, w_policy, b_policy, w_baseline, b_baseline)
    ctx.save_for_backward(core, w_policy, w_baseline)
    return logits, baseline

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g_logits, g_baseline):
    core, w_policy, w_baseline = ctx.saved_tensors
    K = core.shape[-1]
    core2 = core.reshape(-1, K)
    gl = g_logits.reshape(core2.shape[0], -1) if g_logits is not None else None
    gb = g_baseline.reshape(core2.shape[0], 1) if g_baseline is not None else None
    d_core = d_wp = d_bp = d_wv = d_bv = None
    if gl is not None:
      d_wp, d_bp = core2.t() @ gl, gl.sum(0)
      d_core = gl @ w_policy.t()
    if gb is not None:
      d_wv, d_bv = (core2.t() @ gb).reshape(w_baseline.shape), gb.sum(0)
      part = gb @ w_baseline.reshape(1, K)
      d_core = part if d_core is None else d_core + part
    if d_core is not None:
      d_core = d_core.reshape(core.shape)
    return d_core, d_wp, d_bp, d_wv, d_bv


def pack_head_weights(w_policy, b_policy, w_baseline, b_baseline):
  """Packs the heads' parameters for `vtr_policy_head_f32` (once per update)."""
  _ops.require_cuda()
  dev = w_policy.device
  K, A = w_policy.shape
  f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
  w_policy, b_policy = f32(w_policy), f32(b_policy).reshape(A)
  w_baseline, b_baseline = f32(w_baseline).reshape(K), f32(b_baseline).reshape(1)
  lib = _lib.load()
  nbytes = lib.vtr_policy_head_packed_bytes(K, A)
  packed = torch.empty(nbytes // 4, dtype=torch.float32, device=dev)
  with torch.cuda.device(dev):
    rc = lib.vtr_policy_head_prepare_f32(
        w_policy.data_ptr(), b_policy.data_ptr(), w_baseline.data_ptr(), b_baseline.data_ptr(),
        K, A, packed.data_ptr(), _ops._stream(dev))
  _lib.check(rc, 'vtr_policy_head_prepare_f32')
  return packed


def policy_head_forward(core_output, w_policy, b_policy, w_baseline, b_baseline, packed=None):
  """(policy_logits [..., A], baseline [...]) from core_output [..., K]; no autograd."""
  _ops.require_cuda()
  K, A = w_policy.shape
  if core_output.shape[-1] != K:
    raise ValueError('core_output must end with K = %d, got shape %s' % (K, tuple(core_output.shape)))
  dev = core_output.device
  core = core_output.detach()
  if core.dtype != torch.float32 or not core.is_contiguous():
    core = core.to(torch.float32).contiguous()
  if packed is None:
    packed = pack_head_weights(w_policy, b_policy, w_baseline, b_baseline)
  lead = tuple(core.shape[:-1])
  rows = core.numel() // K
  logits = torch.empty(lead + (A,), dtype=torch.float32, device=dev)
  baseline = torch.empty(lead, dtype=torch.float32, device=dev)
  with torch.cuda.device(dev):
    rc = _lib.load().vtr_policy_head_f32(
        core.data_ptr(), packed.data_ptr(), rows, K, A, logits.data_ptr(), baseline.data_ptr(),
        _ops._stream(dev))
  _lib.check(rc, 'vtr_policy_head_f32')
  return logits, baseline


def policy_head(core_output, w_policy, b_policy, w_baseline, b_baseline):
  """experiment.py:199-203 on a B200: returns (policy_logits, baseline),
  differentiable w.r.t. the core output and the four parameters."""
  return _HeadFn.apply(core_output, w_policy, b_policy, w_baseline, b_baseline)
