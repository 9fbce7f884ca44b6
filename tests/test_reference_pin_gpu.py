"""GPU parity against what the reference's own TensorFlow source computes.

The expected side is ``tests/golden/losses_golden.npz`` - ``vtrace.py`` and
``experiment.py:324-343, 366-407`` executed over ``oracle/tf_shim.py`` (see
``tests/golden/make_golden_losses.py``); no code of this repository produced
it.  The CUDA path is called through the reference-shaped Python API, on every
kernel that takes the shape.  Tolerance: 1e-5 * max(1, |ref|) (north_star).
"""
import pytest
import torch

from tests.conftest import Golden, assert_close
from tests.test_reference_pin import (FROM_LOGITS_CASES, HELPER_CASES, LEARNER_CASES, LOSSES_PATH,
                                      shifted)

pytestmark = pytest.mark.gpu

TOL = 1e-5
SPECIAL_A = (3, 4, 6, 9, 18)  # the item kernel keeps rows in registers: compiled widths only


@pytest.fixture(scope='module')
def losses():
  return Golden(LOSSES_PATH)


@pytest.fixture(scope='module')
def api():
  assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
  from scalable_agent_b200 import experiment, vtrace
  return vtrace, experiment


def _dev(x):
  return torch.as_tensor(x).cuda()


def _np(x):
  return x.detach().cpu().numpy()


def _paths(T, B, A):
  paths = ['auto', 'column']
  if B % 4 == 0:
    paths.append('sm_tile')
  if T <= 256 and A in SPECIAL_A:
    paths.append('item')
  if T <= 100 and A <= 40:
    paths.append('chunk')
  return paths


@pytest.mark.parametrize('case', LEARNER_CASES)
def test_learner_block_vs_reference_source(api, losses, case):
  """compute_learner_loss on [T+1] tensors = experiment.py:366-407, forward and backward."""
  _, experiment = api
  c = losses.case(case)
  T1, B, A = c['in/target_policy_logits'].shape
  for path in _paths(T1 - 1, B, A):
    logits = _dev(c['in/target_policy_logits']).requires_grad_(True)
    baseline = _dev(c['in/baseline']).requires_grad_(True)
    out = experiment.compute_learner_loss(
        _dev(c['in/behaviour_policy_logits']), _dev(c['in/actions']), _dev(c['in/rewards']),
        _dev(c['in/done']), logits, baseline, discounting=0.99,
        reward_clipping=str(c['in/reward_clipping']), baseline_cost=0.5, entropy_cost=0.00025,
        return_aux=True, path=path)
    what = '%s path=%s ' % (case, path)
    vt = out.vtrace_returns
    for field in ('vs', 'pg_advantages', 'log_rhos', 'behaviour_action_log_probs', 'target_action_log_probs'):
      assert_close(_np(getattr(vt, field)), c[field], TOL, what + field)
    assert_close(_np(out.policy_gradient_loss), c['pg_loss'], TOL, what + 'pg_loss')
    assert_close(_np(out.baseline_loss), c['baseline_loss'], TOL, what + 'baseline_loss')
    assert_close(_np(out.entropy_loss), c['entropy_loss'], TOL, what + 'entropy_loss')
    assert_close(_np(out.total_loss), c['total_loss'], TOL, what + 'total_loss')
    out.total_loss.backward()
    assert_close(_np(logits.grad), c['d_policy_logits'], TOL, what + 'd_policy_logits')
    assert_close(_np(baseline.grad), c['d_baseline'], TOL, what + 'd_baseline')


@pytest.mark.parametrize('case', LEARNER_CASES)
def test_composed_reference_style_call_vs_reference_source(api, losses, case):
  """The way the reference itself writes it: from_logits, then the three helpers, then autograd."""
  vtrace, experiment = api
  c = losses.case(case)
  s = shifted(c)
  mode = s['reward_clipping']
  rewards = _dev(s['rewards'])
  if mode == 'abs_one':
    clipped = torch.clamp(rewards, -1, 1)
  else:
    squeezed = torch.tanh(rewards / 5.0)
    clipped = torch.where(rewards < 0, .3 * squeezed, squeezed) * 5.
  discounts = (~_dev(s['done'])).float() * 0.99
  logits = _dev(s['target_policy_logits']).requires_grad_(True)
  values = _dev(s['values']).requires_grad_(True)
  vt = vtrace.from_logits(
      behaviour_policy_logits=_dev(s['behaviour_policy_logits']), target_policy_logits=logits,
      actions=_dev(s['actions']), discounts=discounts, rewards=clipped, values=values,
      bootstrap_value=_dev(s['bootstrap_value']))
  total = experiment.compute_policy_gradient_loss(logits, _dev(s['actions']), vt.pg_advantages)
  total = total + 0.5 * experiment.compute_baseline_loss(vt.vs - values)
  total = total + 0.00025 * experiment.compute_entropy_loss(logits)
  assert_close(_np(vt.vs), c['vs'], TOL, 'vs')
  assert_close(_np(vt.pg_advantages), c['pg_advantages'], TOL, 'pg_advantages')
  assert_close(_np(total), c['total_loss'], TOL, 'total_loss')
  total.backward()
  assert_close(_np(logits.grad), c['d_policy_logits'][:-1], TOL, 'd_policy_logits')
  assert_close(_np(values.grad), c['d_baseline'][:-1], TOL, 'd_baseline')


@pytest.mark.parametrize('case', FROM_LOGITS_CASES)
def test_from_logits_vs_reference_source(api, losses, case):
  vtrace, _ = api
  c = losses.case(case)
  got = vtrace.from_logits(
      _dev(c['in/behaviour_policy_logits']), _dev(c['in/target_policy_logits']), _dev(c['in/actions']),
      _dev(c['in/discounts']), _dev(c['in/rewards']), _dev(c['in/values']), _dev(c['in/bootstrap_value']),
      clip_rho_threshold=Golden.threshold(c['in/clip_rho_threshold']),
      clip_pg_rho_threshold=Golden.threshold(c['in/clip_pg_rho_threshold']))
  for field in got._fields:
    assert_close(_np(getattr(got, field)), c[field], TOL, field)


def test_importance_weights_with_trailing_dimensions_vs_reference_source(api, losses):
  vtrace, _ = api
  c = losses.case('importance_weights_trailing')
  got = vtrace.from_importance_weights(
      _dev(c['in/log_rhos']), _dev(c['in/discounts']), _dev(c['in/rewards']), _dev(c['in/values']),
      _dev(c['in/bootstrap_value']))
  assert_close(_np(got.vs), c['vs'], TOL, 'vs')
  assert_close(_np(got.pg_advantages), c['pg_advantages'], TOL, 'pg_advantages')


def test_importance_weights_without_a_batch_dimension_vs_reference_source(api, losses):
  vtrace, _ = api
  c = losses.case('importance_weights_rank1')
  got = vtrace.from_importance_weights(
      _dev(c['in/log_rhos']), _dev(c['in/discounts']), _dev(c['in/rewards']), _dev(c['in/values']),
      _dev(c['in/bootstrap_value']), clip_rho_threshold=3.7, clip_pg_rho_threshold=2.2)
  assert tuple(got.vs.shape) == tuple(c['vs'].shape)
  assert_close(_np(got.vs), c['vs'], TOL, 'vs')
  assert_close(_np(got.pg_advantages), c['pg_advantages'], TOL, 'pg_advantages')


@pytest.mark.parametrize('case', HELPER_CASES)
def test_loss_helpers_vs_reference_source(api, losses, case):
  """experiment.py:324-343, each helper on its own, values and gradients."""
  vtrace, experiment = api
  c = losses.case(case)
  actions = _dev(c['in/actions'])

  adv = _dev(c['in/advantages']).requires_grad_(True)
  bl = experiment.compute_baseline_loss(adv)
  assert_close(_np(bl), c['baseline_loss'], TOL, 'baseline_loss')
  bl.backward()
  assert_close(_np(adv.grad), c['baseline_loss_d_advantages'], TOL, 'd baseline_loss')

  logits = _dev(c['in/logits']).requires_grad_(True)
  ent = experiment.compute_entropy_loss(logits)
  assert_close(_np(ent), c['entropy_loss'], TOL, 'entropy_loss')
  ent.backward()
  assert_close(_np(logits.grad), c['entropy_loss_d_logits'], TOL, 'd entropy_loss')

  logits = _dev(c['in/logits']).requires_grad_(True)
  adv = _dev(c['in/advantages']).requires_grad_(True)
  pg = experiment.compute_policy_gradient_loss(logits, actions, adv)
  assert_close(_np(pg), c['pg_loss'], TOL, 'pg_loss')
  pg.backward()
  assert_close(_np(logits.grad), c['pg_loss_d_logits'], TOL, 'd pg_loss')
  assert adv.grad is None or float(adv.grad.abs().max()) == 0.0  # experiment.py:341

  lp = vtrace.log_probs_from_logits_and_actions(_dev(c['in/logits']), actions)
  assert_close(_np(lp), c['log_probs'], TOL, 'log_probs')
