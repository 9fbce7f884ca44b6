"""The allocation-free entry point, vtr_from_logits_fused_noalloc_f32.

It exists for callers that cannot give the library a workspace because they
cannot allocate device memory themselves: an op embedded in a framework that
passes it inputs and pre-declared outputs only, or a learner in a process
whose GPU memory was handed out at start-up and where cudaMalloc is off limits
afterwards (it may fail, and it synchronises the device under the inference
streams sharing the GPU).  The loss partials then live in an array compiled
into the library.

Checked here, through the C ABI, on two calls in a row with different batch
sizes (so the second call meets the first call's leftovers in the shared
array): every output equals what the ordinary entry point produces with a
caller workspace, to the tolerance of the other GPU tests (1e-5 scaled; in
practice the bits are the same).  Sizes above the built-in limit are refused.

This file sorts last on purpose, so a `pytest -x` run has been through every
other test before it gets here; it also passes when run alone.
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import vtrace_oracle as oracle
from tests.conftest import assert_close

pytestmark = pytest.mark.gpu

KEYS = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
        'discounts', 'rewards', 'values', 'bootstrap_value')
OUT_TB = ('vs', 'pg_advantages', 'log_rhos', 'behaviour_action_log_probs',
          'target_action_log_probs', 'd_values')
NOALLOC_MAX_B = 262144   # VTR_NOALLOC_MAX_B in include/vtrace_b200.h


def _params(_lib, dev_in, T, B, A):
  """A vtr_fused_params over fresh output buffers; returns (params, outputs)."""
  f32 = dict(dtype=torch.float32, device='cuda')
  out = {k: torch.zeros(T, B, **f32) for k in OUT_TB}
  out['d_target_logits'] = torch.zeros(T, B, A, **f32)
  out['losses'] = torch.zeros(4, **f32)
  out['losses_f64'] = torch.zeros(4, dtype=torch.float64, device='cuda')
  p = _lib.FusedParams()
  p.T, p.B, p.A = T, B, A
  for k in KEYS:
    setattr(p, k, dev_in[k].data_ptr())
  p.clip_rho_threshold = p.clip_pg_rho_threshold = 1.0
  p.baseline_cost, p.entropy_cost = 0.5, 0.00025
  for k, v in out.items():
    setattr(p, k, v.data_ptr())
  return p, out


def test_noalloc_entry_point_matches_workspace_entry_point_twice_in_a_row():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

  # two calls in a row; the second is smaller, so stale partials of the first
  # sit right behind the slots it uses
  for call, (T, B, A, seed) in enumerate([(40, 1000, 18, 31), (25, 77, 9, 32)]):
    inp = oracle.synthetic_inputs(T, B, A, seed=seed)
    dev_in = {k: torch.from_numpy(inp[k]).cuda() for k in KEYS}

    p_ref, out_ref = _params(_lib, dev_in, T, B, A)
    nbytes = lib.vtr_fused_workspace_bytes(T, B, A)
    ws = torch.empty(nbytes // 8, dtype=torch.float64, device='cuda')
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p_ref), ws.data_ptr(),
                                       nbytes, stream)
    assert rc == 0, _lib.last_error()

    p_new, out_new = _params(_lib, dev_in, T, B, A)
    rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p_new), stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()

    for k in out_ref:
      assert_close(out_new[k].cpu().numpy(), out_ref[k].cpu().numpy(), 1e-5,
                   'call %d: %s' % (call, k))
    assert torch.equal(out_new['losses_f64'], out_ref['losses_f64'])
    # and both agree with the oracle, so "equal" is not "equally wrong"
    ref = oracle.fused_vtrace_loss(*[inp[k] for k in KEYS])
    assert_close(out_new['losses'].cpu().numpy(),
                 [ref.pg_loss, ref.baseline_loss, ref.entropy_loss,
                  ref.total_loss], 1e-5, 'call %d: losses vs oracle' % call)
    assert_close(out_new['vs'].cpu().numpy(), ref.vs, 1e-5, 'vs vs oracle')


def _free_device_bytes():
  free, _ = torch.cuda.mem_get_info()
  return free


@pytest.mark.parametrize('T,B,A,path,seed', [
    (36, 64 * 148 + 64, 18, 0, 41),   # per-SM kernel: cursor and tickets, partials per granule
    (36, 64 * 148 + 64, 18, 0, 42),   # again: whatever the call before left in the array
    (34, 70000, 6, 0, 43),            # too wide for the per-SM partials: 32-column tiles take over
    (100, 2048, 9, 0, 44),            # cluster kernel, one partial per CTA
    (12, 96, 18, 0, 45),              # small again, behind all that
])
def test_noalloc_entry_point_on_every_kernel_and_allocates_nothing(T, B, A, path, seed):
  """Each kernel the automatic choice can take keeps its partials (and, for
  the per-SM kernel, nothing else) in the built-in array; results match the
  workspace entry point's per-column outputs bit for bit and its losses to
  rounding, and the call leaves the allocator and the device's free memory
  exactly as it found them."""
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  inp = oracle.synthetic_inputs(T, B, A, seed=seed)
  dev_in = {k: torch.from_numpy(inp[k]).cuda() for k in KEYS}
  p_ref, out_ref = _params(_lib, dev_in, T, B, A)
  nbytes = lib.vtr_fused_workspace_bytes(T, B, A)
  ws = torch.empty(nbytes // 8, dtype=torch.float64, device='cuda')
  rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p_ref), ws.data_ptr(), nbytes, stream)
  assert rc == 0, _lib.last_error()
  p_new, out_new = _params(_lib, dev_in, T, B, A)
  p_new.path = path
  # once unmeasured: the first launch of a kernel makes the driver load its
  # code, which is device memory of the driver's, not an allocation of ours
  rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p_new), stream)
  assert rc == 0, _lib.last_error()
  for v in out_new.values():
    v.zero_()
  torch.cuda.synchronize()
  allocated, reserved, free = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved(),
                               _free_device_bytes())
  rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p_new), stream)
  assert rc == 0, _lib.last_error()
  torch.cuda.synchronize()
  assert torch.cuda.memory_allocated() == allocated
  assert torch.cuda.memory_reserved() == reserved
  assert _free_device_bytes() == free
  for k in OUT_TB + ('d_target_logits',):
    assert torch.equal(out_new[k], out_ref[k]), k
  assert_close(out_new['losses_f64'].cpu().numpy(), out_ref['losses_f64'].cpu().numpy(),
               1e-9, 'losses_f64')
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in KEYS])
  assert_close(out_new['losses'].cpu().numpy(),
               [ref.pg_loss, ref.baseline_loss, ref.entropy_loss, ref.total_loss],
               1e-5, 'losses vs oracle')


def test_noalloc_entry_point_without_losses_and_explicit_paths():
  """No losses asked for: no partials, so nothing to fit - any width the
  kernels take goes through.  An explicit kernel whose partials do not fit the
  array is refused with a message that says so."""
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  T, B, A = 8, 65536, 4
  inp = oracle.synthetic_inputs(T, B, A, seed=51)
  dev_in = {k: torch.from_numpy(inp[k]).cuda() for k in KEYS}
  p, out = _params(_lib, dev_in, T, B, A)
  p.path = _lib.PATHS['sm_tile']
  rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p), stream)
  assert rc == -2, 'B/4 granule partials do not fit the built-in array'
  assert 'built-in scratch' in _lib.last_error() and 'VTR_PATH_AUTO' in _lib.last_error()
  p.losses = None
  p.losses_f64 = None
  rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p), stream)
  assert rc == 0, _lib.last_error()
  torch.cuda.synchronize()
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in KEYS])
  assert_close(out['vs'].cpu().numpy(), ref.vs, 1e-5, 'vs')
  assert_close(out['d_target_logits'].cpu().numpy(), ref.d_target_logits, 1e-5, 'd_logits')


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  p = _lib.FusedParams()
  p.T, p.B, p.A = 4, NOALLOC_MAX_B + 1, 4
  rc = lib.vtr_from_logits_fused_noalloc_f32(ctypes.byref(p), None)
  assert rc == -2          # VTR_ERR_UNSUPPORTED, before any launch
  message = _lib.last_error()
  assert 'VTR_NOALLOC_MAX_B' in message and str(NOALLOC_MAX_B) in message
  assert lib.vtr_from_logits_fused_noalloc_f32(None, None) == -1
