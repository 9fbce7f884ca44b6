"""Pins the CPU oracle (oracle/) before anything is checked against it.

* against the frozen outputs of the reference's own numpy ground truth
  (tests/golden/, produced by code lifted from vtrace_test.py);
* against the hard-coded known-answer values of vtrace_test.py:120-156;
* against the live lifted ground truth when /root/reference is mounted;
* loss helpers / gradients against torch float64 autograd of an independent
  restatement (tests/test_reference_pin.py pins them to the reference's own
  TensorFlow source, executed over oracle/tf_shim.py).
"""
import numpy as np
import pytest
import torch

from oracle import cpu_port
from oracle import lift_reference
from oracle import vtrace_oracle as oracle
from tests.conftest import Golden, assert_close


def test_known_answer_anchors(golden):
  # Values the reference's test computes for its own deterministic inputs
  # (vtrace_test.py:120-156), reproduced in SURVEY.md section 4.
  b1 = golden.case('test_vtrace_B1')
  out = oracle.from_importance_weights(
      b1['log_rhos'], b1['discounts'], b1['rewards'], b1['values'],
      b1['bootstrap_value'], 3.7, 2.2)
  assert abs(out.vs[0, 0] - 0.20018187) < 1e-7
  assert abs(out.pg_advantages[0, 0] - 0.20018185) < 1e-7
  b5 = golden.case('test_vtrace_B5')
  out = oracle.from_importance_weights(
      b5['log_rhos'], b5['discounts'], b5['rewards'], b5['values'],
      b5['bootstrap_value'], 3.7, 2.2)
  np.testing.assert_allclose(out.vs[0, :3],
                             [0.9298156, 0.55643743, 0.81689453], atol=2e-7)
  np.testing.assert_allclose(out.pg_advantages[0, :3],
                             [0.92981558, 0.35643743, 0.41689451], atol=2e-7)


@pytest.mark.parametrize('case', ['test_vtrace_B1', 'test_vtrace_B5'])
def test_from_importance_weights_vs_reference_ground_truth(golden, case):
  c = golden.case(case)
  out = oracle.from_importance_weights(
      c['log_rhos'], c['discounts'], c['rewards'], c['values'],
      c['bootstrap_value'], float(c['clip_rho_threshold']),
      float(c['clip_pg_rho_threshold']))
  # float64-input ground truth: tight.
  assert_close(out.vs, c['vs_f64'], 1e-12, 'vs')
  assert_close(out.pg_advantages, c['pg_advantages_f64'], 1e-12, 'pg')
  # as the reference test runs it (float32 accumulator): its own 1e-6.
  assert_close(out.vs, c['vs'], 1e-6, 'vs f32')
  assert_close(out.pg_advantages, c['pg_advantages'], 1e-6, 'pg f32')


@pytest.mark.parametrize('case', ['test_log_probs_B1', 'test_log_probs_B2'])
def test_log_probs_vs_reference(golden, case):
  c = golden.case(case)
  lp = oracle.log_probs_from_logits_and_actions(c['policy_logits'],
                                                c['actions'])
  assert_close(lp, c['log_probs'], 1e-6, 'log_probs')


@pytest.mark.parametrize('case', [
    'test_vtrace_from_logits_B1', 'test_vtrace_from_logits_B2',
    'synthetic_T100_B32_A9', 'synthetic_T20_B96_A18',
    'synthetic_T100_B40_A18_noclip', 'synthetic_T33_B7_A5_clip'])
def test_from_logits_vs_reference_ground_truth(golden, case):
  c = golden.case(case)
  crho = Golden.threshold(c.get('clip_rho_threshold', np.nan))
  cpg = Golden.threshold(c.get('clip_pg_rho_threshold', np.nan))
  out = oracle.from_logits(
      c['behaviour_policy_logits'], c['target_policy_logits'], c['actions'],
      c['discounts'], c['rewards'], c['values'], c['bootstrap_value'],
      crho, cpg)
  for field in out._fields:
    assert_close(getattr(out, field), c[field], 1e-10, field)


@pytest.mark.skipif(not lift_reference.available(),
                    reason='/root/reference not mounted (GPU box)')
def test_live_lifted_ground_truth_matches_oracle():
  ref = lift_reference.lift()
  inp = oracle.synthetic_inputs(50, 12, 6, seed=11)
  lsm_b = np.log(ref._softmax(inp['behaviour_policy_logits'].astype(np.float64)))
  lsm_t = np.log(ref._softmax(inp['target_policy_logits'].astype(np.float64)))
  idx = inp['actions'][..., None].astype(np.int64)
  log_rhos = (np.take_along_axis(lsm_t, idx, -1) -
              np.take_along_axis(lsm_b, idx, -1))[..., 0]
  for crho, cpg in ((1.0, 1.0), (None, None), (3.7, 2.2)):
    gt = ref._ground_truth_calculation(
        discounts=inp['discounts'].astype(np.float64), log_rhos=log_rhos,
        rewards=inp['rewards'].astype(np.float64),
        values=inp['values'].astype(np.float64),
        bootstrap_value=inp['bootstrap_value'].astype(np.float64),
        clip_rho_threshold=crho, clip_pg_rho_threshold=cpg)
    out = oracle.from_logits(
        inp['behaviour_policy_logits'], inp['target_policy_logits'],
        inp['actions'], inp['discounts'], inp['rewards'], inp['values'],
        inp['bootstrap_value'], crho, cpg)
    assert_close(out.vs, gt.vs, 1e-11, 'vs')
    assert_close(out.pg_advantages, gt.pg_advantages, 1e-11, 'pg')


def test_zero_threshold_is_a_real_clip():
  # vtrace.py:236 tests `is not None`; 0.0 must clip rho to 0 (unlike the
  # test helper's truthiness check, vtrace_test.py:53).
  inp = oracle.synthetic_inputs(6, 3, 4, seed=5)
  lp = oracle.from_logits(
      inp['behaviour_policy_logits'], inp['target_policy_logits'],
      inp['actions'], inp['discounts'], inp['rewards'], inp['values'],
      inp['bootstrap_value'], 0.0, 0.0)
  assert_close(lp.vs, inp['values'], 1e-12, 'vs == values when rho_bar = 0')
  assert np.all(lp.pg_advantages == 0.0)


def test_higher_rank_and_rank_errors():
  # vtrace_test.py:249-272
  T, B, C = 4, 3, 42
  rng = np.random.default_rng(0)
  out = oracle.from_importance_weights(
      rng.standard_normal((T, B, 1)), np.full((T, B, 1), 0.9),
      rng.standard_normal((T, B, C)), rng.standard_normal((T, B, C)),
      rng.standard_normal((B, C)))
  assert out.vs.shape == (T, B, C)
  with pytest.raises(ValueError, match='must have rank 2'):
    oracle.from_importance_weights(
        rng.standard_normal((T, B, 1)), np.full((T, B, 1), 0.9),
        rng.standard_normal((T, B, C)), rng.standard_normal((T, B, C)),
        rng.standard_normal((B,)))


def _torch_f64_total_loss(inp, baseline_cost, entropy_cost, crho, cpg):
  """Independent restatement with torch float64 + autograd."""
  t = lambda x: torch.tensor(np.asarray(x, np.float64))
  target = t(inp['target_policy_logits']).requires_grad_(True)
  values = t(inp['values']).requires_grad_(True)
  behaviour = t(inp['behaviour_policy_logits'])
  actions = torch.tensor(inp['actions'].astype(np.int64))
  lsm_t = torch.log_softmax(target, -1)
  lsm_b = torch.log_softmax(behaviour, -1)
  t_lp = lsm_t.gather(-1, actions[..., None])[..., 0]
  b_lp = lsm_b.gather(-1, actions[..., None])[..., 0]
  vt = oracle.from_importance_weights(
      (t_lp - b_lp).detach().numpy(), inp['discounts'], inp['rewards'],
      inp['values'], inp['bootstrap_value'], crho, cpg)
  vs, pg_adv = t(vt.vs), t(vt.pg_advantages)
  pg_loss = torch.sum(-t_lp * pg_adv)
  bl_loss = 0.5 * torch.sum((vs - values) ** 2)
  ent_loss = -torch.sum(-(torch.exp(lsm_t) * lsm_t).sum(-1))
  total = pg_loss + baseline_cost * bl_loss + entropy_cost * ent_loss
  total.backward()
  return (pg_loss.item(), bl_loss.item(), ent_loss.item(), total.item(),
          target.grad.numpy(), values.grad.numpy())


@pytest.mark.parametrize('T,B,A,crho,cpg', [
    (5, 2, 3, None, None), (20, 9, 18, 1.0, 1.0), (100, 4, 9, 3.7, 2.2)])
def test_losses_and_closed_form_gradients_vs_autograd(T, B, A, crho, cpg):
  inp = oracle.synthetic_inputs(T, B, A, seed=T + B)
  bc, ec = 0.5, 0.0033391318945337044
  out = oracle.fused_vtrace_loss(
      inp['behaviour_policy_logits'], inp['target_policy_logits'],
      inp['actions'], inp['discounts'], inp['rewards'], inp['values'],
      inp['bootstrap_value'], crho, cpg, bc, ec)
  pg, bl, ent, total, d_logits, d_values = _torch_f64_total_loss(
      inp, bc, ec, crho, cpg)
  assert_close(out.pg_loss, pg, 1e-12, 'pg_loss')
  assert_close(out.baseline_loss, bl, 1e-12, 'baseline_loss')
  assert_close(out.entropy_loss, ent, 1e-12, 'entropy_loss')
  assert_close(out.total_loss, total, 1e-12, 'total')
  assert_close(out.d_target_logits, d_logits, 1e-12, 'd_logits')
  assert_close(out.d_values, d_values, 1e-12, 'd_values')


def test_reward_clipping_and_discounts():
  r = np.array([-20.0, -1.5, -0.2, 0.0, 0.3, 1.0, 7.0])
  np.testing.assert_allclose(oracle.clip_rewards(r, 'abs_one'),
                             [-1, -1, -0.2, 0, 0.3, 1, 1])
  soft = oracle.clip_rewards(r, 'soft_asymmetric')
  np.testing.assert_allclose(soft[r >= 0], 5 * np.tanh(r[r >= 0] / 5))
  np.testing.assert_allclose(soft[r < 0], 1.5 * np.tanh(r[r < 0] / 5))
  d = oracle.discounts_from_done(np.array([True, False]), 0.99)
  np.testing.assert_allclose(d, [0.0, 0.99])


def test_cpu_port_agrees_with_oracle():
  # The timed CPU baseline (float32, torch-CPU, autograd) computes the same
  # thing as the float64 oracle, to float32 accuracy.
  inp = oracle.synthetic_inputs(30, 16, 9, seed=3)
  t = lambda k: torch.from_numpy(inp[k])
  losses, d_logits, d_values, vs, pg_adv = cpu_port.learner_loss(
      t('behaviour_policy_logits'), t('target_policy_logits'), t('actions'),
      t('discounts'), t('rewards'), t('values'), t('bootstrap_value'))
  out = oracle.fused_vtrace_loss(
      inp['behaviour_policy_logits'], inp['target_policy_logits'],
      inp['actions'], inp['discounts'], inp['rewards'], inp['values'],
      inp['bootstrap_value'])
  assert_close(vs.numpy(), out.vs, 1e-5, 'vs')
  assert_close(pg_adv.numpy(), out.pg_advantages, 1e-5, 'pg')
  assert_close(d_logits.numpy(), out.d_target_logits, 1e-5, 'd_logits')
  assert_close(d_values.numpy(), out.d_values, 1e-5, 'd_values')
  assert_close(losses.numpy(),
               [out.pg_loss, out.baseline_loss, out.entropy_loss], 1e-5,
               'losses')
