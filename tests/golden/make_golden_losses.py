"""Freeze what the reference's OWN TensorFlow code computes for the path.

Run in the build container (where /root/reference is mounted):

    python tests/golden/make_golden_losses.py

TensorFlow 1.x is not installable here, so the reference's source is executed
over ``oracle/tf_shim.py`` (the ~20 elementary TF names it uses, on torch
float64 CPU tensors with autograd) by ``oracle/lift_reference.lift_tf_path``:

* ``vtrace.py``         from_logits, from_importance_weights,
                        log_probs_from_logits_and_actions, as they stand;
* ``experiment.py:324-343``  the three loss helpers, as they stand;
* ``experiment.py:366-407``  build_learner's bootstrap value, T+1 -> T shifts,
                        reward clipping, discounts, V-trace call and weighted
                        total, as they stand.

Every EXPECTED array below comes out of that code; gradients are torch
autograd through it (``tf.stop_gradient`` = ``detach``), i.e. what
``optimizer.minimize`` (experiment.py:415) hands to the network heads.
Nothing from this repository's oracle or kernels is on the expected side; the
oracle only lends its input generator.

Output: tests/golden/losses_golden.npz  (keys "<case>/<field>").
"""
import collections
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import lift_reference  # noqa: E402
from oracle import tf_shim  # noqa: E402

# experiment.py:101-102 (AgentOutput); environments.py:145-146 gives the
# StepOutput order (reward, info, done, observation) unpacked at experiment.py:372.
AgentOutput = collections.namedtuple('AgentOutput', 'action policy_logits baseline')

LEARNER_CASES = [
    # name, T, B, A, reward_clipping, seed
    ('learner_config1', 5, 2, 3, 'abs_one', 11),
    ('learner_abs_one', 20, 8, 9, 'abs_one', 12),
    ('learner_soft_asymmetric', 20, 8, 9, 'soft_asymmetric', 13),
    ('learner_dmlab30', 100, 4, 18, 'abs_one', 14),
    ('learner_wide', 7, 70, 18, 'soft_asymmetric', 15),
    ('learner_atari', 33, 5, 6, 'abs_one', 16),
    ('learner_single_step', 1, 1, 2, 'abs_one', 17),           # the smallest batch there is
    ('learner_one_column', 9, 1, 4, 'soft_asymmetric', 18),
    ('learner_odd_width', 6, 33, 5, 'abs_one', 19),            # one column past a 32-column tile, A not compiled in
]

FROM_LOGITS_CASES = [
    # name, T, B, A, clip_rho, clip_pg_rho, seed
    ('from_logits_default', 20, 8, 9, 1.0, 1.0, 21),
    ('from_logits_unclipped', 20, 8, 9, None, None, 22),
    ('from_logits_3p7_2p2', 13, 5, 18, 3.7, 2.2, 23),
    ('from_logits_rho_only', 13, 5, 4, 2.0, None, 24),
    ('from_logits_long', 100, 3, 9, 1.0, 1.0, 25),
]


def _np(x):
  return x.detach().cpu().numpy()


def learner_inputs(T, B, A, seed):
  """A [T+1] learner batch: what experiment.py:360-362 has in hand."""
  rng = np.random.default_rng(seed)
  behaviour = rng.standard_normal((T + 1, B, A)).astype(np.float32)
  target = (behaviour + 0.5 * rng.standard_normal((T + 1, B, A))).astype(np.float32)
  return dict(
      behaviour_policy_logits=behaviour,
      target_policy_logits=target,
      actions=rng.integers(0, A, size=(T + 1, B)).astype(np.int32),
      rewards=(2.0 * rng.standard_normal((T + 1, B))).astype(np.float32),
      done=rng.random((T + 1, B)) < 0.05,
      baseline=rng.standard_normal((T + 1, B)).astype(np.float32))


def run_learner(ref, inp, reward_clipping, precision):
  tf_shim.set_precision(precision)
  dt = torch.float64 if precision == 'float64' else torch.float32
  flags = types.SimpleNamespace(reward_clipping=reward_clipping, discounting=0.99,
                                baseline_cost=0.5, entropy_cost=0.00025)  # experiment.py:60-67
  logits = torch.tensor(inp['target_policy_logits'], dtype=dt, requires_grad=True)
  baseline = torch.tensor(inp['baseline'], dtype=dt, requires_grad=True)
  actions = torch.tensor(inp['actions'], dtype=torch.int64)
  agent_outputs = AgentOutput(actions, torch.tensor(inp['behaviour_policy_logits'], dtype=dt),
                              torch.zeros_like(baseline).detach())
  learner_outputs = AgentOutput(actions.clone(), logits, baseline)
  rewards = torch.tensor(inp['rewards'], dtype=dt)
  done = torch.tensor(inp['done'])
  placeholder = torch.zeros(inp['rewards'].shape[0], 1)
  env_outputs = (rewards, placeholder, done, placeholder)
  total, vt, clipped, discounts, shifted_learner, shifted_agent, bootstrap = ref.learner_loss(
      flags, agent_outputs, env_outputs, learner_outputs)
  pg = ref.compute_policy_gradient_loss(shifted_learner.policy_logits, shifted_agent.action,
                                        vt.pg_advantages)
  bl = ref.compute_baseline_loss(vt.vs - shifted_learner.baseline)
  ent = ref.compute_entropy_loss(shifted_learner.policy_logits)
  d_logits, d_baseline = torch.autograd.grad(total, [logits, baseline])
  tf_shim.set_precision('float64')
  return dict(
      total_loss=_np(total), pg_loss=_np(pg), baseline_loss=_np(bl), entropy_loss=_np(ent),
      vs=_np(vt.vs), pg_advantages=_np(vt.pg_advantages), log_rhos=_np(vt.log_rhos),
      behaviour_action_log_probs=_np(vt.behaviour_action_log_probs),
      target_action_log_probs=_np(vt.target_action_log_probs),
      clipped_rewards=_np(clipped), discounts=_np(discounts), bootstrap_value=_np(bootstrap),
      d_policy_logits=_np(d_logits), d_baseline=_np(d_baseline))


def main():
  ref = lift_reference.lift_tf_path(tf_shim)
  numpy_truth = lift_reference.lift()
  out = {'meta/glue_lines': np.asarray(ref.glue_lines)}

  # --- experiment.py:366-407 on [T+1] batches --------------------------------
  for name, T, B, A, clipping, seed in LEARNER_CASES:
    inp = learner_inputs(T, B, A, seed)
    for k, v in inp.items():
      out['%s/in/%s' % (name, k)] = v
    out[name + '/in/reward_clipping'] = np.asarray(clipping)
    for k, v in run_learner(ref, inp, clipping, 'float64').items():
      out['%s/%s' % (name, k)] = v
    # The same source in float32, the way the reference itself computes:
    # shows how far an fp32 evaluation sits from the float64 one.
    f32 = run_learner(ref, inp, clipping, 'float32')
    for k in ('total_loss', 'vs', 'pg_advantages'):
      out['%s/f32/%s' % (name, k)] = f32[k]

  # --- vtrace.py:71-161 from_logits, as it stands ----------------------------
  for name, T, B, A, clip_rho, clip_pg_rho, seed in FROM_LOGITS_CASES:
    rng = np.random.default_rng(seed)
    inp = dict(
        behaviour_policy_logits=rng.standard_normal((T, B, A)).astype(np.float32),
        target_policy_logits=(1.5 * rng.standard_normal((T, B, A))).astype(np.float32),
        actions=rng.integers(0, A, size=(T, B)).astype(np.int32),
        discounts=(0.99 * (rng.random((T, B)) > 0.05)).astype(np.float32),
        rewards=rng.standard_normal((T, B)).astype(np.float32),
        values=rng.standard_normal((T, B)).astype(np.float32),
        bootstrap_value=rng.standard_normal((B,)).astype(np.float32))
    for k, v in inp.items():
      out['%s/in/%s' % (name, k)] = v
    out[name + '/in/clip_rho_threshold'] = np.asarray(np.nan if clip_rho is None else clip_rho)
    out[name + '/in/clip_pg_rho_threshold'] = np.asarray(np.nan if clip_pg_rho is None else clip_pg_rho)
    got = ref.vtrace.from_logits(clip_rho_threshold=clip_rho, clip_pg_rho_threshold=clip_pg_rho, **inp)
    for field in got._fields:
      out['%s/%s' % (name, field)] = _np(getattr(got, field))
    # The shim is itself checked here: the reference's TF code run over it
    # must reproduce the reference's own numpy expectation
    # (_ground_truth_calculation, vtrace_test.py:44-83) on the same log_rhos.
    truth = numpy_truth._ground_truth_calculation(
        log_rhos=_np(got.log_rhos), discounts=inp['discounts'].astype(np.float64),
        rewards=inp['rewards'].astype(np.float64), values=inp['values'].astype(np.float64),
        bootstrap_value=inp['bootstrap_value'].astype(np.float64),
        clip_rho_threshold=clip_rho, clip_pg_rho_threshold=clip_pg_rho)
    for field in ('vs', 'pg_advantages'):
      err = np.abs(getattr(truth, field) - _np(getattr(got, field))).max()
      assert err < 1e-12, (name, field, err)

  # --- vtrace.py:164-280 from_importance_weights with trailing dimensions
  # (vtrace_test.py:249-262 feeds [T,B,1] log_rhos against [T,B,42]) ----------
  rng = np.random.default_rng(31)
  T, B, C = 9, 4, 42
  inp = dict(
      log_rhos=rng.standard_normal((T, B, 1)).astype(np.float32),
      discounts=(0.9 * (rng.random((T, B, 1)) > 0.1)).astype(np.float32),
      rewards=rng.standard_normal((T, B, C)).astype(np.float32),
      values=rng.standard_normal((T, B, C)).astype(np.float32),
      bootstrap_value=rng.standard_normal((B, C)).astype(np.float32))
  name = 'importance_weights_trailing'
  for k, v in inp.items():
    out['%s/in/%s' % (name, k)] = v
  got = ref.vtrace.from_importance_weights(**inp)
  out[name + '/vs'] = _np(got.vs)
  out[name + '/pg_advantages'] = _np(got.pg_advantages)

  # ... and without a batch dimension: [T] tensors and a scalar bootstrap value
  rng = np.random.default_rng(32)
  T = 11
  inp = dict(
      log_rhos=rng.standard_normal((T,)).astype(np.float32),
      discounts=(0.9 * (rng.random((T,)) > 0.1)).astype(np.float32),
      rewards=rng.standard_normal((T,)).astype(np.float32),
      values=rng.standard_normal((T,)).astype(np.float32),
      bootstrap_value=np.float32(rng.standard_normal()))
  name = 'importance_weights_rank1'
  for k, v in inp.items():
    out['%s/in/%s' % (name, k)] = np.asarray(v)
  got = ref.vtrace.from_importance_weights(clip_rho_threshold=3.7, clip_pg_rho_threshold=2.2, **inp)
  out[name + '/vs'] = _np(got.vs)
  out[name + '/pg_advantages'] = _np(got.pg_advantages)

  # --- experiment.py:324-343 the three helpers on their own, with gradients ---
  rng = np.random.default_rng(41)
  for name, T, B, A in (('helpers_small', 5, 2, 3), ('helpers_dmlab', 20, 8, 9), ('helpers_wide', 3, 70, 18)):
    logits_np = (2.0 * rng.standard_normal((T, B, A))).astype(np.float32)
    actions_np = rng.integers(0, A, size=(T, B)).astype(np.int32)
    adv_np = rng.standard_normal((T, B)).astype(np.float32)
    out[name + '/in/logits'] = logits_np
    out[name + '/in/actions'] = actions_np
    out[name + '/in/advantages'] = adv_np
    logits = torch.tensor(logits_np, dtype=torch.float64, requires_grad=True)
    adv = torch.tensor(adv_np, dtype=torch.float64, requires_grad=True)
    actions = torch.tensor(actions_np, dtype=torch.int64)
    bl = ref.compute_baseline_loss(adv)
    out[name + '/baseline_loss'] = _np(bl)
    out[name + '/baseline_loss_d_advantages'] = _np(torch.autograd.grad(bl, adv)[0])
    ent = ref.compute_entropy_loss(logits)
    out[name + '/entropy_loss'] = _np(ent)
    out[name + '/entropy_loss_d_logits'] = _np(torch.autograd.grad(ent, logits)[0])
    pg = ref.compute_policy_gradient_loss(logits, actions, adv)
    out[name + '/pg_loss'] = _np(pg)
    d_logits, d_adv = torch.autograd.grad(pg, [logits, adv], allow_unused=True)
    out[name + '/pg_loss_d_logits'] = _np(d_logits)
    # experiment.py:341 stops the gradient into the advantages.
    assert d_adv is None or not d_adv.abs().max() > 0
    lp = ref.vtrace.log_probs_from_logits_and_actions(logits_np, actions_np)
    out[name + '/log_probs'] = _np(lp)

  path = os.path.join(HERE, 'losses_golden.npz')
  np.savez_compressed(path, **out)
  print('wrote %s: %d arrays, %d bytes' % (path, len(out), os.path.getsize(path)))


if __name__ == '__main__':
  main()
