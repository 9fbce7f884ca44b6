"""GPU parity tests: the CUDA path (through the Python API -> ctypes -> C ABI)
against the float64 oracle and the golden fixtures.

Tolerance (BASELINE.json north_star):
    |x - ref| <= 1e-5 * max(1, |ref|)   elementwise and for the loss scalars,
    bitwise equality between two runs on the same inputs.
"""
import numpy as np
import pytest
import torch

from oracle import vtrace_oracle as oracle
from tests.conftest import Golden, assert_close

pytestmark = pytest.mark.gpu

TOL = 1e-5


@pytest.fixture(scope='module')
def api():
  assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
  from scalable_agent_b200 import experiment, vtrace
  return vtrace, experiment


def _dev(x):
  return torch.as_tensor(x).cuda()


def _np(x):
  return x.detach().cpu().numpy()


FUSED_KEYS = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
              'discounts', 'rewards', 'values', 'bootstrap_value')


def _run_fused(experiment, inp, requires_grad=True, **kw):
  args = {k: _dev(inp[k]) for k in FUSED_KEYS}
  if requires_grad:
    args['target_policy_logits'].requires_grad_(True)
    args['values'].requires_grad_(True)
  out = experiment.compute_vtrace_loss(**args, **kw)
  return args, out


def _check_fused(experiment, inp, crho=1.0, cpg=1.0, bc=0.5, ec=0.00025,
                 path='auto', tol=TOL, **kw):
  ref = oracle.fused_vtrace_loss(
      *[inp[k] for k in FUSED_KEYS], crho, cpg, bc, ec,
      reward_clipping=kw.get('reward_clipping'), done=kw.get('done_np'),
      discounting=kw.get('discounting', 0.99))
  call_kw = dict(clip_rho_threshold=crho, clip_pg_rho_threshold=cpg,
                 baseline_cost=bc, entropy_cost=ec, path=path)
  if 'reward_clipping' in kw:
    call_kw['reward_clipping'] = kw['reward_clipping']
  if 'done_np' in kw:
    call_kw['done'] = _dev(kw['done_np'])
    call_kw['discounting'] = kw.get('discounting', 0.99)
  args, out = _run_fused(experiment, inp, **call_kw)
  vt = out.vtrace_returns
  assert_close(_np(vt.vs), ref.vs, tol, 'vs')
  assert_close(_np(vt.pg_advantages), ref.pg_advantages, tol, 'pg_advantages')
  assert_close(_np(vt.log_rhos), ref.log_rhos, tol, 'log_rhos')
  assert_close(_np(vt.behaviour_action_log_probs),
               ref.behaviour_action_log_probs, tol, 'b_lp')
  assert_close(_np(vt.target_action_log_probs), ref.target_action_log_probs,
               tol, 't_lp')
  assert_close(_np(out.policy_gradient_loss), ref.pg_loss, tol, 'pg_loss')
  assert_close(_np(out.baseline_loss), ref.baseline_loss, tol, 'baseline_loss')
  assert_close(_np(out.entropy_loss), ref.entropy_loss, tol, 'entropy_loss')
  assert_close(_np(out.total_loss), ref.total_loss, tol, 'total_loss')
  out.total_loss.backward()
  assert_close(_np(args['target_policy_logits'].grad), ref.d_target_logits,
               tol, 'd_target_logits')
  assert_close(_np(args['values'].grad), ref.d_values, tol, 'd_values')
  return out


# ---------------------------------------------------------------------------
# The reference's own tests (vtrace_test.py), against the frozen expectations
# ---------------------------------------------------------------------------
@pytest.mark.parametrize('case', ['test_log_probs_B1', 'test_log_probs_B2'])
def test_log_probs_from_logits_and_actions(api, golden, case):
  vtrace, _ = api
  c = golden.case(case)
  lp = vtrace.log_probs_from_logits_and_actions(c['policy_logits'], c['actions'])
  assert lp.dtype == torch.float32 and lp.is_cuda
  # the reference's tolerance: assertAllClose default 1e-6 (vtrace_test.py:115)
  np.testing.assert_allclose(_np(lp), c['log_probs'], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize('case', ['test_vtrace_B1', 'test_vtrace_B5'])
def test_vtrace(api, golden, case):
  vtrace, _ = api
  c = golden.case(case)
  out = vtrace.from_importance_weights(
      log_rhos=c['log_rhos'], discounts=c['discounts'], rewards=c['rewards'],
      values=c['values'], bootstrap_value=c['bootstrap_value'],
      clip_rho_threshold=float(c['clip_rho_threshold']),
      clip_pg_rho_threshold=float(c['clip_pg_rho_threshold']))
  assert isinstance(out, vtrace.VTraceReturns)
  assert_close(_np(out.vs), c['vs_f64'], TOL, 'vs')
  assert_close(_np(out.pg_advantages), c['pg_advantages_f64'], TOL, 'pg')
  # and the reference's own float32 expectation at its tolerance (:154-156)
  np.testing.assert_allclose(_np(out.vs), c['vs'], rtol=1e-6, atol=1e-6)
  np.testing.assert_allclose(_np(out.pg_advantages), c['pg_advantages'],
                             rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize('case', ['test_vtrace_from_logits_B1',
                                  'test_vtrace_from_logits_B2'])
def test_vtrace_from_logits(api, golden, case):
  """vtrace_test.py:158-247 (config 1): from_logits == from_importance_weights
  fed the separately computed log-probs; the three extra fields match."""
  vtrace, _ = api
  c = golden.case(case)
  keys = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
          'discounts', 'rewards', 'values', 'bootstrap_value')
  out = vtrace.from_logits(clip_rho_threshold=None, clip_pg_rho_threshold=None,
                           **{k: c[k] for k in keys})
  assert isinstance(out, vtrace.VTraceFromLogitsReturns)
  assert out._fields == ('vs', 'pg_advantages', 'log_rhos',
                         'behaviour_action_log_probs',
                         'target_action_log_probs')
  t_lp = vtrace.log_probs_from_logits_and_actions(c['target_policy_logits'],
                                                  c['actions'])
  b_lp = vtrace.log_probs_from_logits_and_actions(c['behaviour_policy_logits'],
                                                  c['actions'])
  from_iw = vtrace.from_importance_weights(
      log_rhos=t_lp - b_lp, discounts=c['discounts'], rewards=c['rewards'],
      values=c['values'], bootstrap_value=c['bootstrap_value'],
      clip_rho_threshold=None, clip_pg_rho_threshold=None)
  kw = dict(rtol=1e-6, atol=1e-6)
  np.testing.assert_allclose(_np(from_iw.vs), _np(out.vs), **kw)
  np.testing.assert_allclose(_np(from_iw.pg_advantages),
                             _np(out.pg_advantages), **kw)
  np.testing.assert_allclose(_np(b_lp), _np(out.behaviour_action_log_probs), **kw)
  np.testing.assert_allclose(_np(t_lp), _np(out.target_action_log_probs), **kw)
  np.testing.assert_allclose(_np(t_lp - b_lp), _np(out.log_rhos), **kw)
  for f in out._fields:
    assert_close(_np(getattr(out, f)), c[f], TOL, f)
  assert not out.vs.requires_grad and not out.pg_advantages.requires_grad


def test_higher_rank_inputs_for_importance_weights(api):
  """vtrace_test.py:249-259."""
  vtrace, _ = api
  T, B, C = 6, 5, 42
  rng = np.random.default_rng(0)
  log_rhos = rng.standard_normal((T, B, 1)).astype(np.float32)
  discounts = np.full((T, B, 1), 0.9, np.float32)
  rewards = rng.standard_normal((T, B, C)).astype(np.float32)
  values = rng.standard_normal((T, B, C)).astype(np.float32)
  bootstrap = rng.standard_normal((B, C)).astype(np.float32)
  out = vtrace.from_importance_weights(log_rhos, discounts, rewards, values,
                                       bootstrap)
  assert out.vs.shape[-1] == 42 and tuple(out.vs.shape) == (T, B, C)
  ref = oracle.from_importance_weights(log_rhos, discounts, rewards, values,
                                       bootstrap)
  assert_close(_np(out.vs), ref.vs, TOL, 'vs')
  assert_close(_np(out.pg_advantages), ref.pg_advantages, TOL, 'pg')
  # full-width log_rhos/discounts too
  log_rhos_c = np.broadcast_to(log_rhos, (T, B, C)).copy()
  out2 = vtrace.from_importance_weights(log_rhos_c, discounts, rewards, values,
                                        bootstrap, None, None)
  ref2 = oracle.from_importance_weights(log_rhos_c, discounts, rewards, values,
                                        bootstrap, None, None)
  assert_close(_np(out2.vs), ref2.vs, TOL, 'vs')
  assert_close(_np(out2.pg_advantages), ref2.pg_advantages, TOL, 'pg')


def test_inconsistent_rank_inputs_for_importance_weights(api):
  """vtrace_test.py:261-272."""
  vtrace, _ = api
  z = lambda *s: np.zeros(s, np.float32)
  with pytest.raises(ValueError, match='must have rank 2'):
    vtrace.from_importance_weights(z(4, 3, 1), z(4, 3, 1), z(4, 3, 42),
                                   z(4, 3, 42), z(3))


# ---------------------------------------------------------------------------
# Fused forward+backward against the oracle
# ---------------------------------------------------------------------------
@pytest.mark.parametrize('case', [
    'synthetic_T100_B32_A9', 'synthetic_T20_B96_A18',
    'synthetic_T100_B40_A18_noclip', 'synthetic_T33_B7_A5_clip'])
def test_from_logits_vs_reference_ground_truth(api, golden, case):
  vtrace, _ = api
  c = golden.case(case)
  crho = Golden.threshold(c['clip_rho_threshold'])
  cpg = Golden.threshold(c['clip_pg_rho_threshold'])
  out = vtrace.from_logits(*[c[k] for k in FUSED_KEYS], crho, cpg)
  for f in out._fields:
    assert_close(_np(getattr(out, f)), c[f], TOL, f)


@pytest.mark.parametrize('T,B,A', [
    (5, 2, 3),        # config 1 (plumbing)
    (100, 32, 9),     # config 2 (DMLab single level)
    (20, 512, 18),    # short unroll
    (1, 1, 1), (1, 33, 2), (2, 31, 4), (3, 64, 6),
    (7, 65, 5), (40, 37, 7), (9, 100, 32), (6, 34, 100), (4, 3, 256),
    (50, 130, 9), (100, 70, 18),
])
@pytest.mark.parametrize('path', ['column', 'auto'])
def test_fused_matches_oracle(api, T, B, A, path):
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T * 1000 + B + A)
  _check_fused(experiment, inp, path=path)


@pytest.mark.parametrize('share_pct', [60, 93])
@pytest.mark.parametrize('T,B,A', [(12, 148 * 64, 6), (6, 70000, 3), (40, 148 * 160, 18)])
def test_sm_tile_kernel_covers_the_batch_when_shares_fall_short(api, monkeypatch, T, B, A,
                                                                 share_pct):
  """K1c sizes tiles from per-SM speeds on the assumption that every SM runs
  the same number of tiles; when that does not hold the shares need not add
  up to B.  Shrinking every share (test knob) forces it: the tile that draws
  the last ticket must pick up every column nobody was given."""
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T + B + A)
  # the shrunk call comes first: a column it skipped shows whatever the
  # allocator left there, not an earlier identical call's correct value
  monkeypatch.setenv('VTR_SM_SHARE_PCT', str(share_pct))
  _check_fused(experiment, inp, path='sm_tile')
  monkeypatch.delenv('VTR_SM_SHARE_PCT')
  _check_fused(experiment, inp, path='sm_tile')


@pytest.mark.parametrize('T,B,A', [
    (5, 4, 3), (100, 32, 9), (20, 512, 18), (1, 4, 1), (2, 8, 4), (3, 64, 6),
    (7, 68, 5), (40, 36, 7), (9, 100, 32), (6, 36, 100), (4, 4, 256),
    (50, 132, 9), (100, 1000, 18), (100, 4736, 18), (30, 9000, 9),
    (17, 38000, 18), (4, 70000, 3), (3, 140000, 4),
])
def test_sm_tile_kernel_matches_oracle(api, T, B, A):
  """K1c explicitly (B % 4 == 0): every grid shape of its launch heuristics --
  fewer tiles than SMs, one tile per SM with 1..8 column groups, two waves."""
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T * 1000 + B + A)
  _check_fused(experiment, inp, path='sm_tile')


@pytest.mark.parametrize('T,B,A', [
    (5, 2, 3), (100, 32, 9), (20, 512, 18), (1, 1, 3), (33, 7, 4), (64, 65, 6),
    (256, 40, 9), (100, 2048, 9), (50, 130, 18),
])
def test_item_kernel_matches_oracle(api, T, B, A):
  """K2 explicitly: one thread per (t, b) item, T up to 256, compiled-in A."""
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T * 1000 + B + A)
  _check_fused(experiment, inp, path='item')


@pytest.mark.parametrize('T,B,A', [
    (5, 2, 3),          # config 1: one chunk, two lanes
    (100, 32, 9),       # config 2: one cluster of 8 chunks
    (100, 2048, 9),     # config 3
    (20, 16384, 18),    # config 4: two chunks
    (100, 4096, 18),    # the mid point of the sweep
    (1, 1, 1), (2, 33, 2), (13, 31, 4), (14, 64, 6), (25, 65, 5), (40, 37, 7),
    (97, 100, 32), (100, 70, 18), (104, 130, 9), (12, 34, 100), (3, 3, 256),
    (64, 1000, 18), (50, 4100, 3),
])
def test_chunk_kernel_matches_oracle(api, T, B, A):
  """K5 explicitly: T split over a thread-block cluster, chunk aggregates of
  the affine recurrence composed in a fixed order.  Covers one chunk, the
  portable cluster sizes, ragged last chunks and column groups, run-time A
  (rows left in shared memory) and rows that bulk copies cannot take."""
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T * 1000 + B + A)
  _check_fused(experiment, inp, path='chunk')


def test_chunk_kernel_long_unroll_and_limits(api):
  """More steps than 8 chunks hold need a cluster above the portable size:
  either the device schedules it and the result is right, or the call says
  so; the automatic choice never fails."""
  _, experiment = api
  inp = oracle.synthetic_inputs(200, 64, 9, seed=200)
  try:
    _check_fused(experiment, inp, path='chunk')
  except ValueError as e:
    assert 'chunk kernel' in str(e)
  _check_fused(experiment, inp, path='auto')
  with pytest.raises(ValueError, match='chunk kernel'):
    _run_fused(experiment, oracle.synthetic_inputs(300, 4, 9, seed=1), path='chunk')


def test_chunk_kernel_options(api, monkeypatch):
  """Thresholds, reward clipping and done flags on K5; chunk lengths other
  than the default; and the two-launch form (K3 instead of the last CTA)."""
  _, experiment = api
  inp = oracle.synthetic_inputs(24, 200, 9, seed=61, reward_clipping=None)
  inp['rewards'] = inp['raw_rewards']
  _check_fused(experiment, inp, crho=1.0, cpg=1.0, reward_clipping='soft_asymmetric',
               done_np=inp['done'], discounting=0.97, path='chunk',
               ec=0.0033391318945337044)
  _check_fused(experiment, inp, crho=None, cpg=None, reward_clipping='abs_one',
               path='chunk', tol=5e-5)
  for steps in ('3', '8', '13'):
    monkeypatch.setenv('VTR_CHUNK_STEPS', steps)
    _check_fused(experiment, inp, reward_clipping='abs_one', path='chunk')
  monkeypatch.delenv('VTR_CHUNK_STEPS')
  monkeypatch.setenv('VTR_NO_INKERNEL_FINISH', '1')
  _check_fused(experiment, inp, reward_clipping='abs_one', path='chunk')
  _check_fused(experiment, inp, reward_clipping='abs_one', path='item')


def test_item_kernel_limits(api):
  _, experiment = api
  with pytest.raises(ValueError, match='item kernel'):
    _run_fused(experiment, oracle.synthetic_inputs(257, 4, 9, seed=1), path='item')
  with pytest.raises(ValueError, match='item kernel'):
    _run_fused(experiment, oracle.synthetic_inputs(5, 4, 5, seed=1), path='item')


def test_kernels_agree_bitwise(api):
  """The kernels that walk a column step by step (K1a/K1b, K1c) give a column
  the same bits.  K2 and K5 re-associate the recurrence (a scan; chunk
  aggregates), so their vs / pg_advantages / gradients agree to rounding, and
  everything computed before the recurrence is still bit-identical."""
  _, experiment = api
  inp = oracle.synthetic_inputs(60, 2048, 18, seed=77)
  outs = {}
  for path in ('column', 'sm_tile', 'item', 'chunk'):
    args, out = _run_fused(experiment, inp, path=path)
    out.total_loss.backward()
    outs[path] = (out, args['target_policy_logits'].grad, args['values'].grad)
  a = outs['column']
  b = outs['sm_tile']
  for f in a[0].vtrace_returns._fields:
    assert torch.equal(getattr(a[0].vtrace_returns, f), getattr(b[0].vtrace_returns, f)), f
  assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])
  for other in ('item', 'chunk'):
    b = outs[other]
    for f in ('log_rhos', 'behaviour_action_log_probs', 'target_action_log_probs'):
      assert torch.equal(getattr(a[0].vtrace_returns, f), getattr(b[0].vtrace_returns, f)), (other, f)
    for f in ('vs', 'pg_advantages'):
      assert_close(_np(getattr(b[0].vtrace_returns, f)), _np(getattr(a[0].vtrace_returns, f)),
                   2e-6, '%s %s' % (other, f))
    assert_close(_np(b[1]), _np(a[1]), 2e-6, other + ' d_logits')
    assert_close(_np(b[2]), _np(a[2]), 2e-6, other + ' d_values')


def test_sm_tile_refuses_unaligned_batch(api):
  _, experiment = api
  inp = oracle.synthetic_inputs(5, 6, 3, seed=1)
  with pytest.raises(ValueError, match='VTR_PATH_SM_TILE'):
    _run_fused(experiment, inp, path='sm_tile')


@pytest.mark.parametrize('crho,cpg', [(None, None), (3.7, 2.2), (1.0, None),
                                      (0.0, 0.0)])
def test_fused_thresholds(api, crho, cpg):
  _, experiment = api
  inp = oracle.synthetic_inputs(30, 48, 9, seed=9)
  _check_fused(experiment, inp, crho=crho, cpg=cpg, ec=0.0033391318945337044)


def test_fused_reward_clipping_and_done(api):
  """experiment.py:377-384 fused into the prologue."""
  _, experiment = api
  inp = oracle.synthetic_inputs(25, 66, 9, seed=21, reward_clipping=None)
  inp['rewards'] = inp['raw_rewards']
  for mode in ('abs_one', 'soft_asymmetric'):
    _check_fused(experiment, inp, reward_clipping=mode, done_np=inp['done'],
                 discounting=0.97)
  inp = oracle.synthetic_inputs(25, 68, 9, seed=22, reward_clipping=None)
  inp['rewards'] = inp['raw_rewards']
  for path in ('column', 'sm_tile'):
    _check_fused(experiment, inp, reward_clipping='soft_asymmetric',
                 done_np=inp['done'], discounting=0.97, path=path)


def test_config3_dmlab30_large(api):
  """BASELINE.json config 3: T=100, B=2048, A=9, soft_asymmetric."""
  _, experiment = api
  inp = oracle.synthetic_inputs(100, 2048, 9, seed=3,
                                reward_clipping='soft_asymmetric')
  _check_fused(experiment, inp, ec=0.0033391318945337044)


def test_config4_short_unroll(api):
  """BASELINE.json config 4: T=20, B=16384, A=18."""
  _, experiment = api
  inp = oracle.synthetic_inputs(20, 16384, 18, seed=4)
  _check_fused(experiment, inp)


def test_target_shape_full_size(api):
  """The headline shape T=100, B=32768, A=18 against the oracle, plus the
  size-independent properties: run-to-run bitwise determinism and column
  independence (a B-slice computed alone gives the same bits)."""
  _, experiment = api
  T, B, A = 100, 32768, 18
  inp = oracle.synthetic_inputs(T, B, A, seed=1)
  out = _check_fused(experiment, inp)
  args, out2 = _run_fused(experiment, inp)
  assert torch.equal(out.vtrace_returns.vs, out2.vtrace_returns.vs)
  assert torch.equal(out.total_loss, out2.total_loss)
  lo, hi = 4096, 4096 + 640
  sl = {k: (inp[k][:, lo:hi] if inp[k].ndim >= 2 else inp[k][lo:hi]).copy()
        for k in FUSED_KEYS}
  _, out_sl = _run_fused(experiment, sl, path='column')
  assert torch.equal(out_sl.vtrace_returns.vs, out.vtrace_returns.vs[:, lo:hi])
  assert torch.equal(out_sl.vtrace_returns.pg_advantages,
                     out.vtrace_returns.pg_advantages[:, lo:hi])


@pytest.mark.parametrize('path,B', [('auto', 1000), ('chunk', 1000), ('item', 96),
                                    ('column', 1000), ('sm_tile', 1000)])
def test_bitwise_determinism(api, path, B):
  _, experiment = api
  inp = oracle.synthetic_inputs(100, B, 18, seed=5)
  results = []
  for _ in range(3):
    args, out = _run_fused(experiment, inp, path=path)
    out.total_loss.backward()
    results.append((out, args['target_policy_logits'].grad.clone(),
                    args['values'].grad.clone()))
  for out, dl, dv in results[1:]:
    assert torch.equal(out.total_loss, results[0][0].total_loss)
    assert torch.equal(out.policy_gradient_loss, results[0][0].policy_gradient_loss)
    assert torch.equal(out.vtrace_returns.vs, results[0][0].vtrace_returns.vs)
    assert torch.equal(dl, results[0][1]) and torch.equal(dv, results[0][2])


def test_unaligned_views_take_the_generic_loader(api):
  """Rows that are not 16-byte aligned (odd B*A, offset base pointers)."""
  _, experiment = api
  T, B, A = 12, 35, 9            # B*A = 315: rows not 16-byte multiples
  inp = oracle.synthetic_inputs(T, B, A, seed=8)
  _check_fused(experiment, inp)
  # misaligned base pointer: a contiguous tensor carved at a 4-byte offset
  T, B, A = 10, 64, 18
  inp = oracle.synthetic_inputs(T, B, A, seed=12)
  from scalable_agent_b200 import vtrace
  def shifted(x):
    flat = torch.empty(x.size + 1, dtype=torch.float32, device='cuda')
    view = flat[1:].view(x.shape)
    view.copy_(torch.from_numpy(x))
    return view
  out = vtrace.from_logits(
      shifted(inp['behaviour_policy_logits']),
      shifted(inp['target_policy_logits']), inp['actions'], inp['discounts'],
      inp['rewards'], inp['values'], inp['bootstrap_value'])
  ref = oracle.from_logits(*[inp[k] for k in FUSED_KEYS])
  for f in out._fields:
    assert_close(_np(getattr(out, f)), getattr(ref, f), TOL, f)


def test_upstream_gradient_scaling(api):
  _, experiment = api
  inp = oracle.synthetic_inputs(10, 40, 9, seed=2)
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in FUSED_KEYS])
  args, out = _run_fused(experiment, inp)
  (out.total_loss * 0.25).backward()
  assert_close(_np(args['target_policy_logits'].grad),
               0.25 * ref.d_target_logits, TOL, 'scaled d_logits')
  assert_close(_np(args['values'].grad), 0.25 * ref.d_values, TOL,
               'scaled d_values')


# ---------------------------------------------------------------------------
# The stand-alone helpers (experiment.py:324-343) and their composition
# ---------------------------------------------------------------------------
@pytest.mark.parametrize('T,B,A', [(5, 2, 3), (100, 32, 9), (20, 300, 18),
                                   (3, 5, 130)])
def test_loss_helpers_compose_like_the_reference(api, T, B, A):
  """experiment.py:389-407 written with the individual drop-in functions."""
  vtrace, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T + A)
  bc, ec = 0.5, 0.01
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in FUSED_KEYS], 1.0, 1.0, bc, ec)
  logits = _dev(inp['target_policy_logits']).requires_grad_(True)
  baseline = _dev(inp['values']).requires_grad_(True)
  actions = _dev(inp['actions'])
  vtrace_returns = vtrace.from_logits(
      behaviour_policy_logits=_dev(inp['behaviour_policy_logits']),
      target_policy_logits=logits, actions=actions,
      discounts=_dev(inp['discounts']), rewards=_dev(inp['rewards']),
      values=baseline, bootstrap_value=_dev(inp['bootstrap_value']))
  pg_loss = experiment.compute_policy_gradient_loss(
      logits, actions, vtrace_returns.pg_advantages)
  bl_loss = experiment.compute_baseline_loss(vtrace_returns.vs - baseline)
  ent_loss = experiment.compute_entropy_loss(logits)
  total = pg_loss + bc * bl_loss + ec * ent_loss
  assert_close(_np(pg_loss), ref.pg_loss, TOL, 'pg_loss')
  assert_close(_np(bl_loss), ref.baseline_loss, TOL, 'baseline_loss')
  assert_close(_np(ent_loss), ref.entropy_loss, TOL, 'entropy_loss')
  total.backward()
  assert_close(_np(logits.grad), ref.d_target_logits, TOL, 'd_logits')
  assert_close(_np(baseline.grad), ref.d_values, TOL, 'd_values')


def test_log_probs_and_aux_outputs_are_differentiable(api):
  """vtrace.py:156-161: log_rhos and the log-prob fields are not stop_gradient."""
  vtrace, _ = api
  inp = oracle.synthetic_inputs(6, 10, 7, seed=6)
  t64 = torch.tensor(inp['target_policy_logits'], dtype=torch.float64,
                     requires_grad=True)
  b64 = torch.tensor(inp['behaviour_policy_logits'], dtype=torch.float64,
                     requires_grad=True)
  idx = torch.tensor(inp['actions'].astype(np.int64))[..., None]
  w = torch.tensor(np.random.default_rng(0).standard_normal((6, 10)))
  t_lp = torch.log_softmax(t64, -1).gather(-1, idx)[..., 0]
  b_lp = torch.log_softmax(b64, -1).gather(-1, idx)[..., 0]
  ((t_lp - b_lp) * w + 2 * t_lp - 3 * b_lp).sum().backward()

  target = _dev(inp['target_policy_logits']).requires_grad_(True)
  behaviour = _dev(inp['behaviour_policy_logits']).requires_grad_(True)
  out = vtrace.from_logits(behaviour, target, inp['actions'], inp['discounts'],
                           inp['rewards'], inp['values'], inp['bootstrap_value'])
  wd = w.float().cuda()
  (out.log_rhos * wd + 2 * out.target_action_log_probs
   - 3 * out.behaviour_action_log_probs).sum().backward()
  assert_close(_np(target.grad), t64.grad.numpy(), TOL, 'd target')
  assert_close(_np(behaviour.grad), b64.grad.numpy(), TOL, 'd behaviour')

  logits = _dev(inp['target_policy_logits']).requires_grad_(True)
  lp = vtrace.log_probs_from_logits_and_actions(logits, inp['actions'])
  (lp * wd).sum().backward()
  t64.grad = None
  (torch.log_softmax(t64, -1).gather(-1, idx)[..., 0] * w).sum().backward()
  assert_close(_np(logits.grad), t64.grad.numpy(), TOL, 'd log_probs')


def test_c_abi_direct_call(api):
  """Straight through ctypes, no torch op layer: raw pointers in, codes out."""
  import ctypes
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  T, B, A = 8, 50, 9
  inp = oracle.synthetic_inputs(T, B, A, seed=13)
  d = {k: _dev(inp[k]) for k in FUSED_KEYS}
  vs = torch.empty(T, B, device='cuda')
  pg = torch.empty(T, B, device='cuda')
  losses = torch.empty(4, device='cuda')
  losses64 = torch.empty(4, device='cuda', dtype=torch.float64)
  p = _lib.FusedParams()
  p.T, p.B, p.A = T, B, A
  for k in FUSED_KEYS:
    setattr(p, k, d[k].data_ptr())
  p.clip_rho_threshold = p.clip_pg_rho_threshold = 1.0
  p.baseline_cost, p.entropy_cost = 0.5, 0.00025
  p.vs, p.pg_advantages = vs.data_ptr(), pg.data_ptr()
  p.losses, p.losses_f64 = losses.data_ptr(), losses64.data_ptr()
  nbytes = lib.vtr_fused_workspace_bytes(T, B, A)
  ws = torch.empty(nbytes // 8, dtype=torch.float64, device='cuda')
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  # too-small workspace is refused before any launch
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), ws.data_ptr(), 8, stream) == -3
  rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), ws.data_ptr(), nbytes, stream)
  assert rc == 0, _lib.last_error()
  torch.cuda.synchronize()
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in FUSED_KEYS])
  assert_close(_np(vs), ref.vs, TOL, 'vs')
  assert_close(_np(pg), ref.pg_advantages, TOL, 'pg')
  want = [ref.pg_loss, ref.baseline_loss, ref.entropy_loss, ref.total_loss]
  assert_close(_np(losses), want, TOL, 'losses')
  assert_close(_np(losses64), want, TOL, 'losses_f64')


@pytest.mark.parametrize('T,B,A,clip', [(20, 48, 9, 'abs_one'),
                                        (100, 32, 9, 'soft_asymmetric'),
                                        (7, 130, 18, 'abs_one')])
def test_learner_loss_on_t_plus_1_tensors(api, T, B, A, clip):
  """experiment.py:362-407 fed the way build_learner is: [T+1, ...] tensors."""
  _, experiment = api
  rng = np.random.default_rng(T + B)
  learner_logits = rng.standard_normal((T + 1, B, A)).astype(np.float32)
  behaviour_logits = (learner_logits + 0.5 * rng.standard_normal(
      (T + 1, B, A))).astype(np.float32)
  baseline = rng.standard_normal((T + 1, B)).astype(np.float32)
  actions = rng.integers(0, A, (T + 1, B)).astype(np.int32)
  rewards = (2 * rng.standard_normal((T + 1, B))).astype(np.float32)
  done = rng.random((T + 1, B)) < 0.05
  bc, ec, gamma = 0.5, 0.0033391318945337044, 0.99
  ref = oracle.fused_vtrace_loss(
      behaviour_logits[1:], learner_logits[:-1], actions[1:], None,
      rewards[1:], baseline[:-1], baseline[-1], 1.0, 1.0, bc, ec,
      reward_clipping=clip, done=done[1:], discounting=gamma)
  logits_t = _dev(learner_logits).requires_grad_(True)
  baseline_t = _dev(baseline).requires_grad_(True)
  out = experiment.compute_learner_loss(
      _dev(behaviour_logits), _dev(actions), _dev(rewards), _dev(done),
      logits_t, baseline_t, discounting=gamma, reward_clipping=clip,
      baseline_cost=bc, entropy_cost=ec, return_aux=True)
  assert_close(_np(out.vtrace_returns.vs), ref.vs, TOL, 'vs')
  assert_close(_np(out.vtrace_returns.pg_advantages), ref.pg_advantages, TOL, 'pg')
  assert_close(_np(out.vtrace_returns.log_rhos), ref.log_rhos, TOL, 'log_rhos')
  assert_close(_np(out.total_loss), ref.total_loss, TOL, 'total')
  out.total_loss.backward()
  assert tuple(logits_t.grad.shape) == (T + 1, B, A)
  assert_close(_np(logits_t.grad[:-1]), ref.d_target_logits, TOL, 'd_logits')
  assert_close(_np(baseline_t.grad[:-1]), ref.d_values, TOL, 'd_baseline')
  assert float(logits_t.grad[-1].abs().max()) == 0.0
  assert float(baseline_t.grad[-1].abs().max()) == 0.0


def test_many_tiles_and_graceful_fallback(api):
  """Grids far beyond one wave on the 32-column-tile path, and a shape whose
  ring does not fit the per-SM kernel (A=100, 88 columns per SM): the
  automatic choice must fall back to the 32-column tiles, not fail."""
  _, experiment = api
  _check_fused(experiment, oracle.synthetic_inputs(4, 70000, 3, seed=70), path='column')
  _check_fused(experiment, oracle.synthetic_inputs(6, 9600, 100, seed=71), path='auto')


def test_from_logits_wide_batch_without_losses(api):
  """vtrace.from_logits alone (no losses, no gradients, no workspace) on a
  batch wide enough for the per-SM kernel, which then deals tiles evenly."""
  vtrace, _ = api
  inp = oracle.synthetic_inputs(40, 16384, 9, seed=40)
  out = vtrace.from_logits(*[inp[k] for k in FUSED_KEYS])
  ref = oracle.from_logits(*[inp[k] for k in FUSED_KEYS])
  for f in out._fields:
    assert_close(_np(getattr(out, f)), getattr(ref, f), TOL, f)


def test_done_and_clipping_on_the_item_kernel(api):
  _, experiment = api
  inp = oracle.synthetic_inputs(25, 68, 9, seed=23, reward_clipping=None)
  inp['rewards'] = inp['raw_rewards']
  _check_fused(experiment, inp, reward_clipping='abs_one', done_np=inp['done'],
               discounting=0.95, path='item')


# ---------------------------------------------------------------------------
# Losses and gradients without an autograd node
# ---------------------------------------------------------------------------
@pytest.mark.parametrize('T,B,A,clip', [(5, 2, 3, None), (100, 32, 9, 'abs_one'), (20, 300, 18, 'soft_asymmetric')])
def test_loss_and_grads_is_the_autograd_path_without_the_node(api, T, B, A, clip):
  """compute_vtrace_loss_and_grads == compute_vtrace_loss + backward, bit for bit."""
  _, experiment = api
  inp = oracle.synthetic_inputs(T, B, A, seed=T + B, reward_clipping=None)
  kw = dict(reward_clipping=clip, baseline_cost=0.5, entropy_cost=0.01)
  args, out = _run_fused(experiment, inp, **kw)
  out.total_loss.backward()
  plain = {k: _dev(inp[k]) for k in FUSED_KEYS}
  got = experiment.compute_vtrace_loss_and_grads(**plain, **kw)
  assert not got.total_loss.requires_grad and got.d_policy_logits.grad_fn is None
  for a, b in ((got.total_loss, out.total_loss), (got.policy_gradient_loss, out.policy_gradient_loss),
               (got.baseline_loss, out.baseline_loss), (got.entropy_loss, out.entropy_loss),
               (got.vs, out.vtrace_returns.vs), (got.pg_advantages, out.vtrace_returns.pg_advantages),
               (got.d_policy_logits, args['target_policy_logits'].grad), (got.d_baseline, args['values'].grad)):
    assert torch.equal(a.detach(), b.detach())
  # and against the oracle
  ref = oracle.fused_vtrace_loss(*[inp[k] for k in FUSED_KEYS], 1.0, 1.0, 0.5, 0.01, reward_clipping=clip)
  assert_close(_np(got.total_loss), ref.total_loss, TOL, 'total')
  assert_close(_np(got.d_policy_logits), ref.d_target_logits, TOL, 'd_logits')
  assert_close(_np(got.d_baseline), ref.d_values, TOL, 'd_values')
  # host data takes the generic route to the same numbers
  host = experiment.compute_vtrace_loss_and_grads(
      **{k: (inp[k] if k != 'target_policy_logits' else _dev(inp[k])) for k in FUSED_KEYS}, **kw)
  assert torch.equal(host.d_policy_logits, got.d_policy_logits)
  assert torch.equal(host.total_loss, got.total_loss)


def test_learner_loss_and_grads_feeds_the_heads_backward(api):
  """The use the call is for: gradients handed to the backward pass of the heads."""
  _, experiment = api
  T, B, A, K = 12, 40, 9, 32
  g = torch.Generator(device='cuda').manual_seed(5)
  core = torch.randn(T + 1, B, K, generator=g, device='cuda')
  w = (torch.randn(K, A + 1, generator=g, device='cuda') / K ** 0.5)
  behaviour = torch.randn(T + 1, B, A, generator=g, device='cuda')
  actions = torch.randint(0, A, (T + 1, B), generator=g, device='cuda', dtype=torch.int32)
  rewards = torch.randn(T + 1, B, generator=g, device='cuda')
  done = torch.rand(T + 1, B, generator=g, device='cuda') < 0.1

  def heads(weight):
    out = core @ weight
    return out[..., :A].contiguous(), out[..., A].contiguous()

  w1 = w.clone().requires_grad_(True)
  logits, baseline = heads(w1)
  experiment.compute_learner_loss(behaviour, actions, rewards, done, logits, baseline).total_loss.backward()

  w2 = w.clone().requires_grad_(True)
  logits, baseline = heads(w2)
  r = experiment.compute_learner_loss_and_grads(behaviour, actions, rewards, done, logits.detach(),
                                                baseline.detach())
  assert tuple(r.d_policy_logits.shape) == (T + 1, B, A) and tuple(r.d_baseline.shape) == (T + 1, B)
  assert float(r.d_policy_logits[-1].abs().max()) == 0.0 and float(r.d_baseline[-1].abs().max()) == 0.0
  torch.autograd.backward([logits, baseline], [r.d_policy_logits, r.d_baseline])
  assert torch.equal(w1.grad, w2.grad)
