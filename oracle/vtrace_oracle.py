"""CPU oracle for the V-trace + IMPALA-loss hot path.  TEST INFRASTRUCTURE ONLY.

This module is the checker, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  The shipped path is the CUDA library under
``scalable_agent_b200/csrc`` and fails loudly when that library is missing.

It is a float64 numpy restatement, O(T) in the unroll length, of

* ``vtrace.py:45-68``   log_probs_from_logits_and_actions
* ``vtrace.py:71-161``  from_logits
* ``vtrace.py:164-280`` from_importance_weights
* ``experiment.py:324-343`` compute_baseline_loss / compute_entropy_loss /
  compute_policy_gradient_loss
* ``experiment.py:377-384`` reward clipping and discounts
* ``experiment.py:401-407`` the weighted total loss

plus the closed-form gradients of that total loss w.r.t. the target logits and
the baseline values (what ``optimizer.minimize`` at ``experiment.py:415``
back-propagates into the network heads).

Pinning.  Two independent sources, both the reference's own code, frozen as
fixtures under ``tests/golden/``:

1. V-trace (``from_importance_weights``, the action log-probs, ``from_logits``)
   against the reference's numpy ground truth ``_ground_truth_calculation``
   (``vtrace_test.py:44-83``), lifted by ``oracle/lift_reference.lift`` ->
   ``vtrace_golden.npz`` (``tests/golden/make_golden.py``).
2. Everything - V-trace as ``vtrace.py`` itself writes it, the three loss
   helpers (``experiment.py:324-343``), bootstrap value, T+1 -> T shifts,
   reward clipping, discounts and the weighted total (``:366-407``), and the
   gradients into the policy logits and the baseline - against the
   reference's TensorFlow source EXECUTED over ``oracle/tf_shim.py``, a
   stand-in for the ~20 elementary TF names that code uses, on torch float64
   with autograd (``oracle/lift_reference.lift_tf_path`` ->
   ``losses_golden.npz``, ``tests/golden/make_golden_losses.py``).  The
   generator checks the stand-in itself: the TF code run over it reproduces
   source 1 to 1e-12.  ``tests/test_reference_pin.py`` holds this oracle to
   those fixtures at 1e-12.

What stays restated rather than executed: the TF primitives behind the
stand-in (``reduce_sum``, ``softmax``, ``sparse_softmax_cross_entropy`` ...),
one line each from the TF 1.x API documentation, because TensorFlow 1.x cannot
be installed here.  RMSProp (``oracle/rmsprop_oracle.py``) is TF library code
that is not in the reference tree: **parity unpinned**.
"""
import collections

import numpy as np

VTraceFromLogitsReturns = collections.namedtuple(
    'VTraceFromLogitsReturns',
    ['vs', 'pg_advantages', 'log_rhos',
     'behaviour_action_log_probs', 'target_action_log_probs'])

VTraceReturns = collections.namedtuple('VTraceReturns', 'vs pg_advantages')

FusedReturns = collections.namedtuple(
    'FusedReturns',
    ['vs', 'pg_advantages', 'log_rhos', 'behaviour_action_log_probs',
     'target_action_log_probs', 'pg_loss', 'baseline_loss', 'entropy_loss',
     'total_loss', 'd_target_logits', 'd_values'])


def _f64(x):
  return np.asarray(x, dtype=np.float64)


def log_softmax(logits):
  """Stable log-softmax over the last axis (tf.nn.log_softmax)."""
  logits = _f64(logits)
  m = np.max(logits, axis=-1, keepdims=True)
  z = logits - m
  return z - np.log(np.sum(np.exp(z), axis=-1, keepdims=True))


def log_probs_from_logits_and_actions(policy_logits, actions):
  """vtrace.py:45-68: -sparse_softmax_cross_entropy_with_logits."""
  policy_logits = _f64(policy_logits)
  actions = np.asarray(actions, dtype=np.int64)
  if policy_logits.ndim != 3:
    raise ValueError('policy_logits must have rank 3')
  if actions.ndim != 2:
    raise ValueError('actions must have rank 2')
  lsm = log_softmax(policy_logits)
  return np.take_along_axis(lsm, actions[..., None], axis=-1)[..., 0]


def from_importance_weights(log_rhos, discounts, rewards, values,
                            bootstrap_value, clip_rho_threshold=1.0,
                            clip_pg_rho_threshold=1.0):
  """vtrace.py:164-280, float64, one reverse pass over T.

  Thresholds follow the reference's ``is not None`` test (vtrace.py:215,236,
  270): 0.0 is a real clip, unlike vtrace_test.py:53,56 where it is falsy.
  Extra trailing dimensions broadcast as numpy/TF do (vtrace.py:178-180).
  """
  log_rhos = _f64(log_rhos)
  discounts = _f64(discounts)
  rewards = _f64(rewards)
  values = _f64(values)
  bootstrap_value = _f64(bootstrap_value)
  rho_rank = log_rhos.ndim
  for name, arr, rank in (('values', values, rho_rank),
                          ('bootstrap_value', bootstrap_value, rho_rank - 1),
                          ('discounts', discounts, rho_rank),
                          ('rewards', rewards, rho_rank)):
    if arr.ndim != rank:
      raise ValueError('%s must have rank %d' % (name, rank))

  rhos = np.exp(log_rhos)                                    # :235
  if clip_rho_threshold is not None:                         # :236-239
    clipped_rhos = np.minimum(float(clip_rho_threshold), rhos)
  else:
    clipped_rhos = rhos
  cs = np.minimum(1.0, rhos)                                 # :241
  values_t_plus_1 = np.concatenate(                          # :243-244
      [values[1:], bootstrap_value[None]], axis=0)
  deltas = clipped_rhos * (rewards + discounts * values_t_plus_1 - values)

  seq_len = values.shape[0]
  acc = np.zeros_like(bootstrap_value + deltas[0])           # :256, broadcast
  vs_minus_v_xs = np.empty((seq_len,) + acc.shape, dtype=np.float64)
  for t in range(seq_len - 1, -1, -1):                       # :250-262
    acc = deltas[t] + discounts[t] * cs[t] * acc
    vs_minus_v_xs[t] = acc
  vs = vs_minus_v_xs + values                                # :265
  vs_t_plus_1 = np.concatenate([vs[1:], bootstrap_value[None]], axis=0)
  if clip_pg_rho_threshold is not None:                      # :270-274
    clipped_pg_rhos = np.minimum(float(clip_pg_rho_threshold), rhos)
  else:
    clipped_pg_rhos = rhos
  pg_advantages = clipped_pg_rhos * (
      rewards + discounts * vs_t_plus_1 - values)            # :275-276
  return VTraceReturns(vs=vs, pg_advantages=pg_advantages)


def from_logits(behaviour_policy_logits, target_policy_logits, actions,
                discounts, rewards, values, bootstrap_value,
                clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0):
  """vtrace.py:71-161."""
  behaviour_policy_logits = _f64(behaviour_policy_logits)
  target_policy_logits = _f64(target_policy_logits)
  if behaviour_policy_logits.ndim != 3 or target_policy_logits.ndim != 3:
    raise ValueError('policy logits must have rank 3')
  target_lp = log_probs_from_logits_and_actions(target_policy_logits, actions)
  behaviour_lp = log_probs_from_logits_and_actions(
      behaviour_policy_logits, actions)
  log_rhos = target_lp - behaviour_lp                        # :147
  ret = from_importance_weights(
      log_rhos, discounts, rewards, values, bootstrap_value,
      clip_rho_threshold, clip_pg_rho_threshold)
  return VTraceFromLogitsReturns(
      log_rhos=log_rhos, behaviour_action_log_probs=behaviour_lp,
      target_action_log_probs=target_lp, **ret._asdict())


def compute_baseline_loss(advantages):
  """experiment.py:324-328."""
  return 0.5 * np.sum(np.square(_f64(advantages)))


def compute_entropy_loss(logits):
  """experiment.py:331-335."""
  lsm = log_softmax(logits)
  entropy_per_timestep = np.sum(-np.exp(lsm) * lsm, axis=-1)
  return -np.sum(entropy_per_timestep)


def compute_policy_gradient_loss(logits, actions, advantages):
  """experiment.py:338-343."""
  cross_entropy = -log_probs_from_logits_and_actions(logits, actions)
  return np.sum(cross_entropy * _f64(advantages))


def clip_rewards(rewards, mode):
  """experiment.py:377-382.  mode in {None/'none', 'abs_one', 'soft_asymmetric'}."""
  rewards = _f64(rewards)
  if mode in (None, 'none'):
    return rewards
  if mode == 'abs_one':
    return np.clip(rewards, -1.0, 1.0)
  if mode == 'soft_asymmetric':
    squeezed = np.tanh(rewards / 5.0)
    return np.where(rewards < 0, 0.3 * squeezed, squeezed) * 5.0
  raise ValueError('unknown reward clipping mode %r' % (mode,))


def discounts_from_done(done, discounting):
  """experiment.py:384."""
  return (~np.asarray(done, dtype=bool)).astype(np.float64) * float(discounting)


def loss_gradients(target_policy_logits, actions, values, vs, pg_advantages,
                   baseline_cost, entropy_cost):
  """Closed-form d(total_loss)/d(target_logits), d(total_loss)/d(values).

  total_loss = pg + baseline_cost*bl + entropy_cost*ent (experiment.py:401-407)
  with vs and pg_advantages constants (vtrace.py:279-280, experiment.py:341):
    d/dlogits = (p - onehot(a)) * pg_adv + entropy_cost * p * (log p + H)
    d/dvalues = baseline_cost * (values - vs)
  where H = -sum_a p log p.
  """
  lsm = log_softmax(target_policy_logits)
  p = np.exp(lsm)
  onehot = np.zeros_like(p)
  np.put_along_axis(onehot, np.asarray(actions, np.int64)[..., None], 1.0, -1)
  entropy = -np.sum(p * lsm, axis=-1, keepdims=True)
  pg = _f64(pg_advantages)[..., None]
  d_logits = (p - onehot) * pg + float(entropy_cost) * p * (lsm + entropy)
  d_values = float(baseline_cost) * (_f64(values) - _f64(vs))
  return d_logits, d_values


def fused_vtrace_loss(behaviour_policy_logits, target_policy_logits, actions,
                      discounts, rewards, values, bootstrap_value,
                      clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
                      baseline_cost=0.5, entropy_cost=0.00025,
                      reward_clipping=None, done=None, discounting=0.99):
  """The whole path of experiment.py:377-407 after the T+1 -> T shift.

  If ``done`` is given, ``discounts`` is ignored and rebuilt from it
  (experiment.py:384).  Returns every V-trace field, the three losses, the
  total and both gradients, all float64.
  """
  rewards = clip_rewards(rewards, reward_clipping)
  if done is not None:
    discounts = discounts_from_done(done, discounting)
  vt = from_logits(behaviour_policy_logits, target_policy_logits, actions,
                   discounts, rewards, values, bootstrap_value,
                   clip_rho_threshold, clip_pg_rho_threshold)
  values = _f64(values)
  pg_loss = compute_policy_gradient_loss(
      target_policy_logits, actions, vt.pg_advantages)
  baseline_loss = compute_baseline_loss(vt.vs - values)
  entropy_loss = compute_entropy_loss(target_policy_logits)
  total = pg_loss + baseline_cost * baseline_loss + entropy_cost * entropy_loss
  d_logits, d_values = loss_gradients(
      target_policy_logits, actions, values, vt.vs, vt.pg_advantages,
      baseline_cost, entropy_cost)
  return FusedReturns(
      vs=vt.vs, pg_advantages=vt.pg_advantages, log_rhos=vt.log_rhos,
      behaviour_action_log_probs=vt.behaviour_action_log_probs,
      target_action_log_probs=vt.target_action_log_probs,
      pg_loss=pg_loss, baseline_loss=baseline_loss, entropy_loss=entropy_loss,
      total_loss=total, d_target_logits=d_logits, d_values=d_values)


def synthetic_inputs(T, B, A, seed=1, reward_clipping='abs_one',
                     discounting=0.99, done_prob=0.01):
  """Synthetic learner batch of SURVEY.md section 8(d), as float32/int32 numpy.

  behaviour ~ N(0,1); target = behaviour + N(0,0.5^2); actions ~ U{0..A-1};
  done ~ Bernoulli(done_prob); raw rewards ~ N(0,2^2) then clipped per
  ``reward_clipping``; values, bootstrap ~ N(0,1).
  """
  rng = np.random.default_rng(seed)
  behaviour = rng.standard_normal((T, B, A)).astype(np.float32)
  target = (behaviour +
            0.5 * rng.standard_normal((T, B, A))).astype(np.float32)
  actions = rng.integers(0, A, size=(T, B)).astype(np.int32)
  done = rng.random((T, B)) < done_prob
  raw_rewards = (2.0 * rng.standard_normal((T, B))).astype(np.float32)
  rewards = clip_rewards(raw_rewards, reward_clipping).astype(np.float32)
  discounts = discounts_from_done(done, discounting).astype(np.float32)
  values = rng.standard_normal((T, B)).astype(np.float32)
  bootstrap = rng.standard_normal((B,)).astype(np.float32)
  return dict(behaviour_policy_logits=behaviour, target_policy_logits=target,
              actions=actions, discounts=discounts, rewards=rewards,
              values=values, bootstrap_value=bootstrap, done=done,
              raw_rewards=raw_rewards)
