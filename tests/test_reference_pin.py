"""Pins the oracle to the reference's own TensorFlow code for the path.

``tests/golden/losses_golden.npz`` holds what ``vtrace.py`` and the loss block
of ``experiment.py`` (:324-343, :366-407) compute when their source is executed
over ``oracle/tf_shim.py`` (``tests/golden/make_golden_losses.py``): V-trace
fields, reward clipping, discounts, the three losses, the weighted total and
the gradients into the network heads.  Here the float64 oracle must agree with
those to 1e-12; ``tests/test_reference_pin_gpu.py`` holds the kernels to them.
"""
import os

import numpy as np
import pytest

from tests.conftest import ROOT, Golden, assert_close
from oracle import lift_reference
from oracle import vtrace_oracle as oracle

LOSSES_PATH = os.path.join(ROOT, 'tests', 'golden', 'losses_golden.npz')
F64 = 1e-12

LEARNER_CASES = ['learner_config1', 'learner_abs_one', 'learner_soft_asymmetric', 'learner_dmlab30',
                 'learner_wide', 'learner_atari', 'learner_single_step', 'learner_one_column',
                 'learner_odd_width']
FROM_LOGITS_CASES = ['from_logits_default', 'from_logits_unclipped', 'from_logits_3p7_2p2',
                     'from_logits_rho_only', 'from_logits_long']
HELPER_CASES = ['helpers_small', 'helpers_dmlab', 'helpers_wide']


@pytest.fixture(scope='module')
def losses():
  return Golden(LOSSES_PATH)


def shifted(c):
  """experiment.py:366-375: what build_learner feeds V-trace and the losses."""
  return dict(
      behaviour_policy_logits=c['in/behaviour_policy_logits'][1:],
      target_policy_logits=c['in/target_policy_logits'][:-1],
      actions=c['in/actions'][1:],
      rewards=c['in/rewards'][1:],
      done=c['in/done'][1:],
      values=c['in/baseline'][:-1],
      bootstrap_value=c['in/baseline'][-1],
      reward_clipping=str(c['in/reward_clipping']))


def test_fixture_covers_the_loss_block(losses):
  first, last = (int(x) for x in losses.case('meta')['glue_lines'])
  assert (first, last) == (366, 407)
  for name in (LEARNER_CASES + FROM_LOGITS_CASES + HELPER_CASES +
               ['importance_weights_trailing', 'importance_weights_rank1']):
    assert name in losses.cases, name


@pytest.mark.parametrize('case', LEARNER_CASES)
def test_oracle_learner_block_vs_reference_source(losses, case):
  c = losses.case(case)
  s = shifted(c)
  assert_close(oracle.clip_rewards(s['rewards'], s['reward_clipping']), c['clipped_rewards'], F64, 'clip')
  assert_close(oracle.discounts_from_done(s['done'], 0.99), c['discounts'], F64, 'discounts')
  got = oracle.fused_vtrace_loss(
      s['behaviour_policy_logits'], s['target_policy_logits'], s['actions'], None, s['rewards'],
      s['values'], s['bootstrap_value'], 1.0, 1.0, 0.5, 0.00025,
      reward_clipping=s['reward_clipping'], done=s['done'], discounting=0.99)
  for field in ('vs', 'pg_advantages', 'log_rhos', 'behaviour_action_log_probs',
                'target_action_log_probs', 'pg_loss', 'baseline_loss', 'entropy_loss', 'total_loss'):
    assert_close(getattr(got, field), c[field], F64, field)
  # Gradients on the [T+1] tensors: the last step only feeds the bootstrap
  # value, which V-trace stops (vtrace.py:279-280), so its gradient is zero.
  assert_close(got.d_target_logits, c['d_policy_logits'][:-1], F64, 'd_policy_logits')
  assert_close(got.d_values, c['d_baseline'][:-1], F64, 'd_baseline')
  assert not c['d_policy_logits'][-1].any() and not c['d_baseline'][-1].any()


@pytest.mark.parametrize('case', LEARNER_CASES)
def test_float32_evaluation_of_the_reference_source_is_within_1e4(losses, case):
  """SURVEY.md 8(c)'s second leg (<= 1e-4 against an fp32 evaluation), as far
  as it can be had without TensorFlow: the reference source in float32."""
  c = losses.case(case)
  for field in ('total_loss', 'vs', 'pg_advantages'):
    assert_close(c['f32/' + field], c[field], 1e-4, field)


@pytest.mark.parametrize('case', FROM_LOGITS_CASES)
def test_oracle_from_logits_vs_reference_source(losses, case):
  c = losses.case(case)
  got = oracle.from_logits(
      c['in/behaviour_policy_logits'], c['in/target_policy_logits'], c['in/actions'], c['in/discounts'],
      c['in/rewards'], c['in/values'], c['in/bootstrap_value'],
      Golden.threshold(c['in/clip_rho_threshold']), Golden.threshold(c['in/clip_pg_rho_threshold']))
  for field in got._fields:
    assert_close(getattr(got, field), c[field], F64, field)


def test_oracle_importance_weights_with_trailing_dimensions(losses):
  c = losses.case('importance_weights_trailing')
  got = oracle.from_importance_weights(c['in/log_rhos'], c['in/discounts'], c['in/rewards'], c['in/values'],
                                       c['in/bootstrap_value'])
  assert_close(got.vs, c['vs'], F64, 'vs')
  assert_close(got.pg_advantages, c['pg_advantages'], F64, 'pg_advantages')


def test_oracle_importance_weights_without_a_batch_dimension(losses):
  c = losses.case('importance_weights_rank1')
  got = oracle.from_importance_weights(
      c['in/log_rhos'][:, None], c['in/discounts'][:, None], c['in/rewards'][:, None], c['in/values'][:, None],
      c['in/bootstrap_value'][None], 3.7, 2.2)
  assert_close(got.vs[:, 0], c['vs'], F64, 'vs')
  assert_close(got.pg_advantages[:, 0], c['pg_advantages'], F64, 'pg_advantages')


@pytest.mark.parametrize('case', HELPER_CASES)
def test_oracle_loss_helpers_vs_reference_source(losses, case):
  c = losses.case(case)
  logits, actions, adv = c['in/logits'], c['in/actions'], c['in/advantages']
  assert_close(oracle.compute_baseline_loss(adv), c['baseline_loss'], F64, 'baseline')
  assert_close(oracle.compute_entropy_loss(logits), c['entropy_loss'], F64, 'entropy')
  assert_close(oracle.compute_policy_gradient_loss(logits, actions, adv), c['pg_loss'], F64, 'pg')
  assert_close(oracle.log_probs_from_logits_and_actions(logits, actions), c['log_probs'], F64, 'log_probs')
  assert_close(adv.astype(np.float64), c['baseline_loss_d_advantages'], F64, 'd baseline')
  zeros = np.zeros_like(adv)
  # d(pg)/dlogits = (p - onehot) adv;  d(entropy loss)/dlogits = p (log p + H)
  d_pg, _ = oracle.loss_gradients(logits, actions, zeros, zeros, adv, 0.0, 0.0)
  d_ent, _ = oracle.loss_gradients(logits, actions, zeros, zeros, zeros, 0.0, 1.0)
  assert_close(d_pg, c['pg_loss_d_logits'], F64, 'd pg')
  assert_close(d_ent, c['entropy_loss_d_logits'], F64, 'd entropy')


@pytest.mark.skipif(not lift_reference.tf_path_available(), reason='/root/reference not mounted (GPU box)')
def test_fixture_is_what_the_reference_source_computes_today(losses):
  """Re-runs the lifted reference code and compares with the committed file."""
  import sys
  sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
  import make_golden_losses as gen
  from oracle import tf_shim
  ref = lift_reference.lift_tf_path(tf_shim)
  name, T, B, A, clipping, seed = gen.LEARNER_CASES[2]
  live = gen.run_learner(ref, gen.learner_inputs(T, B, A, seed), clipping, 'float64')
  c = losses.case(name)
  for field, value in live.items():
    np.testing.assert_array_equal(value, c[field], err_msg=field)
