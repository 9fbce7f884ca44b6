"""Drop-in for the loss helpers of the reference's `experiment.py`, on a B200.

`compute_baseline_loss`, `compute_entropy_loss` and
`compute_policy_gradient_loss` keep the reference's names, arguments and
values (`/root/reference/experiment.py:324-343`) and are differentiable the
way TF autodiff makes them (`experiment.py:415`).  `compute_vtrace_loss` is the
fused form of the whole block `experiment.py:377-407`: one pass over the
logits yields V-trace, the three losses, their weighted total and the
gradients w.r.t. the learner's logits and baseline.

Only the learner-loss path of `experiment.py` lives here; the TF1 training
loop, agent network and environments are out of scope.
"""
import collections

import torch

from scalable_agent_b200 import _ext
from scalable_agent_b200 import _lib
from scalable_agent_b200 import _ops
from scalable_agent_b200 import vtrace
from scalable_agent_b200.vtrace import _assert_rank, _device_of, _threshold, _to

VTraceLossReturns = collections.namedtuple(
    'VTraceLossReturns',
    ['total_loss', 'policy_gradient_loss', 'baseline_loss', 'entropy_loss',
     'vtrace_returns'])


def _device_ready(*tensors_and_dtypes):
  """True if every (tensor, dtype) pair is already what the kernels take: a
  contiguous CUDA tensor of that dtype.  Then the C++ host layer does the
  rest (checks included); anything else goes through the coercions of the
  Python layer first."""
  for t, dtype in tensors_and_dtypes:
    if not (type(t) is torch.Tensor or isinstance(t, torch.Tensor)):
      return False
    if not t.is_cuda or t.dtype is not dtype or not t.is_contiguous():
      return False
  return True


def _returns(total, losses, vs, pg, log_rhos, b_lp, t_lp):
  return VTraceLossReturns(
      total_loss=total, policy_gradient_loss=losses[0],
      baseline_loss=losses[1], entropy_loss=losses[2],
      vtrace_returns=vtrace.VTraceFromLogitsReturns(
          vs=vs, pg_advantages=pg, log_rhos=log_rhos,
          behaviour_action_log_probs=b_lp, target_action_log_probs=t_lp))


def compute_baseline_loss(advantages):
  """0.5 * sum(advantages**2) (experiment.py:324-328)."""
  dev = _device_of(advantages)
  advantages = _to(advantages, torch.float32, dev)
  return _ops.BaselineLoss.apply(advantages)


def compute_entropy_loss(logits):
  """-sum_t H(softmax(logits_t)) (experiment.py:331-335)."""
  dev = _device_of(logits)
  logits = _to(logits, torch.float32, dev)
  return _ops.EntropyLoss.apply(logits)


def compute_policy_gradient_loss(logits, actions, advantages):
  """sum_t CE(logits_t, a_t) * stop_gradient(advantages_t) (experiment.py:338-343)."""
  dev = _device_of(logits, actions, advantages)
  logits = _to(logits, torch.float32, dev)
  actions = _to(actions, torch.int32, dev)
  advantages = _to(advantages, torch.float32, dev).detach()
  if tuple(actions.shape) != tuple(logits.shape[:-1]):
    raise ValueError('actions must have the shape of logits minus its last dim')
  if tuple(advantages.shape) != tuple(actions.shape):
    raise ValueError('advantages must have the shape of actions')
  return _ops.PolicyGradientLoss.apply(logits, actions, advantages)


def compute_vtrace_loss(
    behaviour_policy_logits, target_policy_logits, actions, discounts,
    rewards, values, bootstrap_value,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
    baseline_cost=.5, entropy_cost=.00025,
    reward_clipping=None, done=None, discounting=.99,
    return_aux=True, path='auto'):
  """Fused learner loss: the block `experiment.py:377-407` in one pass.

  Equivalent to

      vtrace_returns = vtrace.from_logits(...)                       # :389-397
      total  = compute_policy_gradient_loss(target_policy_logits, actions,
                                            vtrace_returns.pg_advantages)
      total += baseline_cost * compute_baseline_loss(vtrace_returns.vs - values)
      total += entropy_cost * compute_entropy_loss(target_policy_logits)  # :401-407

  but reads each input once, and `total_loss.backward()` costs no second pass
  over the logits: the gradients w.r.t. `target_policy_logits` and `values`
  are produced by the same kernel.  Defaults are the reference's flags
  (`experiment.py:70-72`).

  Args:
    behaviour_policy_logits, target_policy_logits: float32 [T, B, NUM_ACTIONS].
    actions: int32 [T, B].
    discounts: float32 [T, B]; ignored (may be None) when `done` is given.
    rewards, values: float32 [T, B].  bootstrap_value: float32 [B].
    reward_clipping: None, 'abs_one' or 'soft_asymmetric' applied to
      `rewards` inside the kernel (experiment.py:377-382).
    done: optional bool [T, B]; discounts = (~done) * discounting (:384).
    return_aux: also materialise log_rhos and the two action log-prob fields.
    path: 'auto', 'column', 'sm_tile', 'item' or 'chunk' (kernel choice; see
      DESIGN.md).

  Returns:
    VTraceLossReturns(total_loss, policy_gradient_loss, baseline_loss,
    entropy_loss, vtrace_returns).  Only `total_loss` is differentiable; the
    component losses are reported unweighted, for logging.
  """
  if reward_clipping not in _lib.REWARD_CLIP_MODES:
    raise ValueError('unknown reward_clipping %r' % (reward_clipping,))
  if path not in _lib.PATHS:
    raise ValueError('unknown path %r' % (path,))
  ext = None if torch.compiler.is_compiling() else _ext.get()
  f32 = torch.float32
  if ext is not None and _device_ready(
      (behaviour_policy_logits, f32), (target_policy_logits, f32),
      (actions, torch.int32), (rewards, f32), (values, f32), (bootstrap_value, f32),
      (done, torch.bool) if done is not None else (discounts, f32)):
    # The C++ host layer: same checks, same kernels, a fraction of the host time.
    return _returns(*ext.fused_loss(
        behaviour_policy_logits, target_policy_logits, actions,
        None if done is not None else discounts, rewards, values, bootstrap_value,
        done.view(torch.uint8) if done is not None else None,
        _threshold(clip_rho_threshold, 'clip_rho_threshold'),
        _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'),
        float(baseline_cost), float(entropy_cost), float(discounting),
        _lib.REWARD_CLIP_MODES[reward_clipping], bool(return_aux), _lib.PATHS[path]))

  dev = _device_of(behaviour_policy_logits, target_policy_logits, actions,
                   discounts, rewards, values, bootstrap_value)
  behaviour_policy_logits = _to(behaviour_policy_logits, torch.float32, dev)
  target_policy_logits = _to(target_policy_logits, torch.float32, dev)
  actions = _to(actions, torch.int32, dev)
  rewards = _to(rewards, torch.float32, dev)
  values = _to(values, torch.float32, dev)
  bootstrap_value = _to(bootstrap_value, torch.float32, dev)
  if done is not None:
    done = _to(done, torch.bool, dev).view(torch.uint8)
    discounts = torch.empty(0, dtype=torch.float32, device=dev)
  elif discounts is None:
    raise ValueError('one of discounts / done must be given')
  else:
    discounts = _to(discounts, torch.float32, dev)
  if reward_clipping not in _lib.REWARD_CLIP_MODES:
    raise ValueError('unknown reward_clipping %r' % (reward_clipping,))
  if path not in _lib.PATHS:
    raise ValueError('unknown path %r' % (path,))
  clip_rho_threshold = _threshold(clip_rho_threshold, 'clip_rho_threshold')
  clip_pg_rho_threshold = _threshold(clip_pg_rho_threshold,
                                     'clip_pg_rho_threshold')

  _assert_rank(behaviour_policy_logits, 3, 'behaviour_policy_logits')
  _assert_rank(target_policy_logits, 3, 'target_policy_logits')
  T, B, A = target_policy_logits.shape
  if tuple(behaviour_policy_logits.shape) != (T, B, A):
    raise ValueError('behaviour and target logits differ in shape')
  per_step = [(actions, 'actions'), (rewards, 'rewards'), (values, 'values')]
  per_step.append((done, 'done') if done is not None
                  else (discounts, 'discounts'))
  for t, n in per_step:
    _assert_rank(t, 2, n)
    if tuple(t.shape) != (T, B):
      raise ValueError('%s must have shape [T, B] = %s' % (n, (T, B)))
  _assert_rank(bootstrap_value, 1, 'bootstrap_value')
  if tuple(bootstrap_value.shape) != (B,):
    raise ValueError('bootstrap_value must have shape [B]')

  total, losses, vs, pg, log_rhos, b_lp, t_lp = _ops.FusedLoss.apply(
      behaviour_policy_logits, target_policy_logits, actions, discounts,
      rewards, values, bootstrap_value, done, clip_rho_threshold,
      clip_pg_rho_threshold, float(baseline_cost), float(entropy_cost),
      float(discounting), _lib.REWARD_CLIP_MODES[reward_clipping],
      bool(return_aux), _lib.PATHS[path])
  aux = (log_rhos, b_lp, t_lp) if return_aux else (None, None, None)
  return VTraceLossReturns(
      total_loss=total, policy_gradient_loss=losses[0],
      baseline_loss=losses[1], entropy_loss=losses[2],
      vtrace_returns=vtrace.VTraceFromLogitsReturns(
          vs=vs, pg_advantages=pg, log_rhos=aux[0],
          behaviour_action_log_probs=aux[1], target_action_log_probs=aux[2]))


def compute_learner_loss(
    behaviour_policy_logits, actions, rewards, done,
    learner_policy_logits, learner_baseline,
    discounting=.99, reward_clipping='abs_one',
    baseline_cost=.5, entropy_cost=.00025,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
    return_aux=False, path='auto'):
  """The loss block of `build_learner` (`experiment.py:362-407`) on its own inputs.

  Every argument has T+1 time steps, as `build_learner` receives them:
  `behaviour_policy_logits` / `actions` are the actors' `agent_outputs`,
  `rewards` / `done` the `env_outputs`, `learner_policy_logits` /
  `learner_baseline` the result of `agent.unroll`.  This function does what the
  reference does between the unroll and the optimiser:

    bootstrap_value = learner_baseline[-1]                       # :366
    agent/env outputs [1:], learner outputs [:-1]                # :372-375
    clipped_rewards = clip(rewards)                              # :377-382
    discounts = float(~done) * discounting                       # :384
    vtrace.from_logits(...) and the three weighted losses        # :389-407

  in one kernel launch: the shifts are views (time is the outermost
  dimension), clipping and discounts happen inside the kernel, and the
  gradients w.r.t. `learner_policy_logits` and `learner_baseline` come back in
  their full [T+1, ...] shape.  Defaults are the reference's flags
  (`experiment.py:70-74`).

  Returns a `VTraceLossReturns`; only `total_loss` is differentiable.
  """
  if reward_clipping not in _lib.REWARD_CLIP_MODES:
    raise ValueError('unknown reward_clipping %r' % (reward_clipping,))
  if path not in _lib.PATHS:
    raise ValueError('unknown path %r' % (path,))
  ext = None if torch.compiler.is_compiling() else _ext.get()
  f32 = torch.float32
  if ext is not None and _device_ready(
      (learner_policy_logits, f32), (learner_baseline, f32), (behaviour_policy_logits, f32),
      (actions, torch.int32), (rewards, f32), (done, torch.bool)):
    return _returns(*ext.fused_learner_loss(
        learner_policy_logits, learner_baseline, behaviour_policy_logits, actions,
        rewards, done.view(torch.uint8),
        _threshold(clip_rho_threshold, 'clip_rho_threshold'),
        _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'),
        float(baseline_cost), float(entropy_cost), float(discounting),
        _lib.REWARD_CLIP_MODES[reward_clipping], bool(return_aux), _lib.PATHS[path]))

  dev = _device_of(learner_policy_logits, learner_baseline,
                   behaviour_policy_logits)
  learner_policy_logits = _to(learner_policy_logits, torch.float32, dev)
  learner_baseline = _to(learner_baseline, torch.float32, dev)
  behaviour_policy_logits = _to(behaviour_policy_logits, torch.float32, dev)
  actions = _to(actions, torch.int32, dev)
  rewards = _to(rewards, torch.float32, dev)
  done = _to(done, torch.bool, dev).view(torch.uint8)
  if reward_clipping not in _lib.REWARD_CLIP_MODES:
    raise ValueError('unknown reward_clipping %r' % (reward_clipping,))
  if path not in _lib.PATHS:
    raise ValueError('unknown path %r' % (path,))
  _assert_rank(learner_policy_logits, 3, 'learner_policy_logits')
  _assert_rank(behaviour_policy_logits, 3, 'behaviour_policy_logits')
  T1, B, A = learner_policy_logits.shape
  if T1 < 2:
    raise ValueError('need at least 2 time steps (T+1), got %d' % T1)
  if tuple(behaviour_policy_logits.shape) != (T1, B, A):
    raise ValueError('behaviour and learner logits differ in shape')
  for t, n in ((learner_baseline, 'learner_baseline'), (actions, 'actions'),
               (rewards, 'rewards'), (done, 'done')):
    _assert_rank(t, 2, n)
    if tuple(t.shape) != (T1, B):
      raise ValueError('%s must have shape [T+1, B] = %s' % (n, (T1, B)))

  total, losses, vs, pg, log_rhos, b_lp, t_lp = _ops.FusedLearnerLoss.apply(
      learner_policy_logits, learner_baseline, behaviour_policy_logits,
      actions, rewards, done,
      _threshold(clip_rho_threshold, 'clip_rho_threshold'),
      _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'),
      float(baseline_cost), float(entropy_cost), float(discounting),
      _lib.REWARD_CLIP_MODES[reward_clipping], bool(return_aux),
      _lib.PATHS[path])
  aux = (log_rhos, b_lp, t_lp) if return_aux else (None, None, None)
  return VTraceLossReturns(
      total_loss=total, policy_gradient_loss=losses[0],
      baseline_loss=losses[1], entropy_loss=losses[2],
      vtrace_returns=vtrace.VTraceFromLogitsReturns(
          vs=vs, pg_advantages=pg, log_rhos=aux[0],
          behaviour_action_log_probs=aux[1], target_action_log_probs=aux[2]))


VTraceLossAndGrads = collections.namedtuple(
    'VTraceLossAndGrads',
    ['total_loss', 'policy_gradient_loss', 'baseline_loss', 'entropy_loss',
     'vs', 'pg_advantages', 'd_policy_logits', 'd_baseline'])


def _loss_and_grads(total, losses, vs, pg, d_logits, d_baseline):
  return VTraceLossAndGrads(
      total_loss=total, policy_gradient_loss=losses[0], baseline_loss=losses[1],
      entropy_loss=losses[2], vs=vs, pg_advantages=pg,
      d_policy_logits=d_logits, d_baseline=d_baseline)


def compute_vtrace_loss_and_grads(
    behaviour_policy_logits, target_policy_logits, actions, discounts,
    rewards, values, bootstrap_value,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
    baseline_cost=.5, entropy_cost=.00025,
    reward_clipping=None, done=None, discounting=.99, path='auto'):
  """`compute_vtrace_loss` + `total_loss.backward()` as one call without an autograd node.

  The fused pass produces the losses and the gradients of the total w.r.t.
  `target_policy_logits` and `values` in its one launch; here they are handed
  back directly, for a learner that continues the backward pass itself:

      r = compute_vtrace_loss_and_grads(...)
      torch.autograd.backward([logits, baseline], [r.d_policy_logits, r.d_baseline])

  which is what `optimizer.minimize(total_loss)` (`experiment.py:415`) does to
  the heads.  Same kernels and numbers as `compute_vtrace_loss`; the host side
  is one C++ call (about 10 us against about 60 us through the autograd
  engine), which is what counts at small shapes.  Nothing returned is
  differentiable.
  """
  ext = None if torch.compiler.is_compiling() else _ext.get()
  f32 = torch.float32
  if (ext is not None and reward_clipping in _lib.REWARD_CLIP_MODES and path in _lib.PATHS
      and _device_ready(
          (behaviour_policy_logits, f32), (target_policy_logits, f32),
          (actions, torch.int32), (rewards, f32), (values, f32), (bootstrap_value, f32),
          (done, torch.bool) if done is not None else (discounts, f32))):
    return _loss_and_grads(*ext.fused_loss_and_grads(
        behaviour_policy_logits, target_policy_logits, actions,
        None if done is not None else discounts, rewards, values, bootstrap_value,
        done.view(torch.uint8) if done is not None else None,
        _threshold(clip_rho_threshold, 'clip_rho_threshold'),
        _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'),
        float(baseline_cost), float(entropy_cost), float(discounting),
        _lib.REWARD_CLIP_MODES[reward_clipping], _lib.PATHS[path]))
  # Anything else (host data, other dtypes, no C++ layer): through the
  # differentiable call on leaves of our own.
  dev = _device_of(behaviour_policy_logits, target_policy_logits, actions,
                   discounts, rewards, values, bootstrap_value)
  logits = _to(target_policy_logits, f32, dev).detach().clone().requires_grad_(True)
  baseline = _to(values, f32, dev).detach().clone().requires_grad_(True)
  out = compute_vtrace_loss(
      behaviour_policy_logits, logits, actions, discounts, rewards, baseline,
      bootstrap_value, clip_rho_threshold, clip_pg_rho_threshold, baseline_cost,
      entropy_cost, reward_clipping, done, discounting, return_aux=False, path=path)
  out.total_loss.backward()
  return VTraceLossAndGrads(
      out.total_loss.detach(), out.policy_gradient_loss, out.baseline_loss, out.entropy_loss,
      out.vtrace_returns.vs, out.vtrace_returns.pg_advantages, logits.grad, baseline.grad)


def compute_learner_loss_and_grads(
    behaviour_policy_logits, actions, rewards, done,
    learner_policy_logits, learner_baseline,
    discounting=.99, reward_clipping='abs_one',
    baseline_cost=.5, entropy_cost=.00025,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0, path='auto'):
  """`compute_learner_loss` + `total_loss.backward()` as one call without an autograd node.

  Arguments as `compute_learner_loss` ([T+1, ...] tensors, `experiment.py:362-407`);
  `d_policy_logits` and `d_baseline` come back [T+1, ...] with a zero last
  step.  See `compute_vtrace_loss_and_grads`.
  """
  ext = None if torch.compiler.is_compiling() else _ext.get()
  f32 = torch.float32
  if (ext is not None and reward_clipping in _lib.REWARD_CLIP_MODES and path in _lib.PATHS
      and _device_ready(
          (learner_policy_logits, f32), (learner_baseline, f32), (behaviour_policy_logits, f32),
          (actions, torch.int32), (rewards, f32), (done, torch.bool))):
    return _loss_and_grads(*ext.fused_learner_loss_and_grads(
        learner_policy_logits, learner_baseline, behaviour_policy_logits, actions,
        rewards, done.view(torch.uint8),
        _threshold(clip_rho_threshold, 'clip_rho_threshold'),
        _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'),
        float(baseline_cost), float(entropy_cost), float(discounting),
        _lib.REWARD_CLIP_MODES[reward_clipping], _lib.PATHS[path]))
  dev = _device_of(learner_policy_logits, learner_baseline, behaviour_policy_logits)
  logits = _to(learner_policy_logits, f32, dev).detach().clone().requires_grad_(True)
  baseline = _to(learner_baseline, f32, dev).detach().clone().requires_grad_(True)
  out = compute_learner_loss(
      behaviour_policy_logits, actions, rewards, done, logits, baseline,
      discounting, reward_clipping, baseline_cost, entropy_cost,
      clip_rho_threshold, clip_pg_rho_threshold, return_aux=False, path=path)
  out.total_loss.backward()
  return VTraceLossAndGrads(
      out.total_loss.detach(), out.policy_gradient_loss, out.baseline_loss, out.entropy_loss,
      out.vtrace_returns.vs, out.vtrace_returns.pg_advantages, logits.grad, baseline.grad)
