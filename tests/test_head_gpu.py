"""The tensor-core heads kernel (vtr_policy_head_f32, csrc/vtrace_head.cu)
against float64: experiment.py:199-203 applied over [T*B, K] (:237).

Tolerance: the same 1e-5 scaled bound as the V-trace outputs that are computed
from these logits - a single TF32 pass would miss it by two orders of
magnitude (checked below), the 3xTF32 products meet it."""
import numpy as np
import pytest
import torch

from oracle import vtrace_oracle as oracle
from tests.conftest import assert_close

pytestmark = pytest.mark.gpu


def _params(K, A, seed):
  rng = np.random.default_rng(seed)
  return ((rng.standard_normal((K, A)) / np.sqrt(K)).astype(np.float32),
          rng.standard_normal(A).astype(np.float32),
          (rng.standard_normal(K) / np.sqrt(K)).astype(np.float32),
          rng.standard_normal(1).astype(np.float32))


@pytest.mark.parametrize('T,B,K,A', [
    (100, 32, 256, 9),      # config 2: 3200 rows = 25 tiles
    (5, 2, 256, 3),         # config 1: one ragged tile
    (20, 300, 256, 18),     # 6000 rows: ragged last tile, several tiles per SM at most
    (100, 2048, 256, 9),    # config 3: 1600 tiles, 10-11 per SM through the ring
    (7, 33, 64, 31),        # fewest k-blocks, widest head
    (3, 50, 32, 1),         # one k-block, one action
])
def test_heads_match_float64(T, B, K, A):
  assert torch.cuda.is_available()
  from scalable_agent_b200 import heads
  rng = np.random.default_rng(T * 31 + B)
  core = rng.standard_normal((T, B, K)).astype(np.float32)
  wp, bp, wv, bv = _params(K, A, seed=K + A)
  logits, baseline = heads.policy_head_forward(
      torch.from_numpy(core).cuda(), torch.from_numpy(wp).cuda(), torch.from_numpy(bp).cuda(),
      torch.from_numpy(wv).cuda(), torch.from_numpy(bv).cuda())
  torch.cuda.synchronize()
  ref_logits = core.astype(np.float64) @ wp.astype(np.float64) + bp.astype(np.float64)
  ref_baseline = core.astype(np.float64) @ wv.astype(np.float64) + float(bv[0])
  assert tuple(logits.shape) == (T, B, A) and tuple(baseline.shape) == (T, B)
  assert_close(logits.cpu().numpy(), ref_logits, 1e-5, 'policy_logits')
  assert_close(baseline.cpu().numpy(), ref_baseline, 1e-5, 'baseline')
  # and it is not a plain TF32 product: that would be ~1e-3 off
  tf32 = lambda x: (x.view(np.uint32) & np.uint32(0xffffe000)).view(np.float32).astype(np.float64)
  one_pass = tf32(core) @ tf32(wp) + bp
  assert np.abs(one_pass - ref_logits).max() > 50 * np.abs(logits.cpu().numpy() - ref_logits).max()


def test_heads_then_fused_loss_match_the_oracle():
  """Heads -> fused V-trace + losses, against the float64 oracle fed float64
  logits: the whole learner block from the core output on."""
  assert torch.cuda.is_available()
  from scalable_agent_b200 import experiment, heads
  T, B, K, A = 40, 96, 256, 18
  inp = oracle.synthetic_inputs(T, B, A, seed=11)
  rng = np.random.default_rng(5)
  core = rng.standard_normal((T, B, K)).astype(np.float32)
  wp, bp, wv, bv = _params(K, A, seed=3)
  ref_logits = core.astype(np.float64) @ wp.astype(np.float64) + bp.astype(np.float64)
  ref_values = core.astype(np.float64) @ wv.astype(np.float64) + float(bv[0])
  ref = oracle.fused_vtrace_loss(inp['behaviour_policy_logits'], ref_logits, inp['actions'],
                                 inp['discounts'], inp['rewards'], ref_values, inp['bootstrap_value'])
  dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
  core_t = dev(core).requires_grad_(True)
  params = [dev(x).requires_grad_(True) for x in (wp, bp, wv, bv)]
  logits, values = heads.policy_head(core_t, *params)
  out = experiment.compute_vtrace_loss(
      dev(inp['behaviour_policy_logits']), logits, dev(inp['actions']), dev(inp['discounts']),
      dev(inp['rewards']), values, dev(inp['bootstrap_value']))
  assert_close(out.vtrace_returns.vs.cpu().numpy(), ref.vs, 1e-5, 'vs')
  assert_close(out.vtrace_returns.pg_advantages.cpu().numpy(), ref.pg_advantages, 1e-5, 'pg')
  assert_close(out.total_loss.detach().cpu().numpy(), ref.total_loss, 1e-5, 'total')
  out.total_loss.backward()
  d_core = ref.d_target_logits @ wp.astype(np.float64).T + ref.d_values[..., None] * wv.astype(np.float64)
  assert_close(core_t.grad.cpu().numpy(), d_core, 2e-5, 'd_core')
  d_wp = core.reshape(-1, K).astype(np.float64).T @ ref.d_target_logits.reshape(-1, A)
  assert_close(params[0].grad.cpu().numpy(), d_wp, 1e-4, 'd_w_policy')   # torch matmul, fp32 sums of 3840 terms


def test_heads_reject_bad_shapes():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  assert lib.vtr_policy_head_packed_bytes(256, 18) == (2 * 32 * 256 + 32) * 4
  assert lib.vtr_policy_head_f32(None, None, 128, 256, 18, None, None, None) == -1
  x = torch.zeros(128 * 256, device='cuda')
  assert lib.vtr_policy_head_f32(x.data_ptr(), x.data_ptr(), 128, 100, 18, x.data_ptr(), x.data_ptr(), None) == -2
  assert 'multiple of 32' in _lib.last_error()
  assert lib.vtr_policy_head_f32(x.data_ptr(), x.data_ptr(), 128, 256, 32, x.data_ptr(), x.data_ptr(), None) == -2
