"""CPU-side checks of the C-ABI boundary: the library loads, exports every
symbol include/vtrace_b200.h declares, and rejects bad arguments before any
launch (no GPU needed: these calls return before touching the device)."""
import ctypes
import os
import re

import pytest

from scalable_agent_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'vtrace_b200.h')


def _declared_symbols():
  text = open(HEADER).read()
  text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
  return sorted(set(re.findall(r'\b(vtr_[a-z0-9_]+)\s*\(', text)))


def test_header_symbols_are_exported_and_bound():
  declared = _declared_symbols()
  assert 'vtr_from_logits_fused_f32' in declared and len(declared) >= 15
  lib = _lib.load()
  for name in declared:
    assert hasattr(lib, name), '%s declared in the header but not exported' % name
  # the Python binding covers exactly the header
  assert sorted(_lib.SIGNATURES) == declared


def test_version_and_sizes():
  lib = _lib.load()
  assert lib.vtr_version() == 200
  assert lib.vtr_fused_workspace_bytes(100, 32768, 18) == 8192 + 8192 * 24
  assert lib.vtr_fused_workspace_bytes(5, 2, 3) == 8192 + 16 * 24  # K5: up to 16 CTAs per 32 columns
  assert lib.vtr_fused_workspace_bytes(5, 700, 3) == 8192 + 700 * 24  # K2: one partial per column
  assert lib.vtr_fused_workspace_bytes(0, 4, 4) == 0
  assert lib.vtr_loss_workspace_bytes(1) == 8
  assert lib.vtr_loss_workspace_bytes(0) == 0


def test_struct_layout_matches_header():
  # 4 ints, 8 pointers, 5 floats + 3 ints, 9 pointers, 1 pointer + 4 ints + 2 pointers on LP64.
  assert ctypes.sizeof(_lib.FusedParams) == 16 + 8 * 8 + 32 + 9 * 8 + 8 + 16 + 8 + 8


def test_argument_errors_are_negative_codes_with_messages():
  lib = _lib.load()
  assert lib.vtr_from_logits_fused_f32(None, None, 0, None) == -1
  assert 'NULL' in _lib.last_error()
  p = _lib.FusedParams()
  p.T, p.B, p.A = 5, 2, 0
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), None, 0, None) == -1
  assert 'positive' in _lib.last_error()
  p.A = _lib.VTR_MAX_ACTIONS + 1
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), None, 0, None) == -2
  p.A = 3
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), None, 0, None) == -1
  with pytest.raises(ValueError, match='required input pointer'):
    _lib.check(-1, 'vtr_from_logits_fused_f32')
  assert lib.vtr_reduce_losses(None, 0, 0.5, 0.1, None, None, None) == -1
  assert lib.vtr_log_probs_from_logits_and_actions_f32(None, None, 4, 3, None, None) == -1
  assert lib.vtr_baseline_loss_f32(None, 0, None, None, 0, None) == -1


def test_ops_fail_loudly_without_cuda():
  import torch
  if torch.cuda.is_available():
    pytest.skip('CUDA present')
  from scalable_agent_b200 import vtrace
  z = torch.zeros
  with pytest.raises(RuntimeError, match='no CPU fallback'):
    vtrace.from_logits(z(2, 2, 3), z(2, 2, 3), z(2, 2), z(2, 2), z(2, 2),
                       z(2, 2), z(2))
