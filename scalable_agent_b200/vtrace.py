"""Drop-in for the reference's `vtrace` module, running on a B200.

Same function names, argument order, keyword names, defaults and namedtuples
as `/root/reference/vtrace.py` (:37-42, :45, :71-75, :164-167).  Tensors are
torch CUDA tensors; array-likes and host tensors are coerced to float32 /
int32 and moved to the current CUDA device, the way the reference coerces
with `tf.convert_to_tensor` (:61-62, :128-132, :210-214).  All arithmetic runs
in the hand-written sm_100a kernels of libvtrace_b200.so; there is no CPU path.
"""
import collections
import math

import torch

from scalable_agent_b200 import _lib
from scalable_agent_b200 import _ops

VTraceFromLogitsReturns = collections.namedtuple(
    'VTraceFromLogitsReturns',
    ['vs', 'pg_advantages', 'log_rhos',
     'behaviour_action_log_probs', 'target_action_log_probs'])

VTraceReturns = collections.namedtuple('VTraceReturns', 'vs pg_advantages')


def _device_of(*xs):
  for x in xs:
    if isinstance(x, torch.Tensor) and x.is_cuda:
      return x.device
  _ops.require_cuda()
  return torch.device('cuda', torch.cuda.current_device())


def _to(x, dtype, device):
  """convert_to_tensor: dtype coercion + move to the device, contiguous."""
  if not isinstance(x, torch.Tensor):
    x = torch.as_tensor(x)
  if x.dtype != dtype or x.device != device:
    x = x.to(device=device, dtype=dtype, non_blocking=True)
  return x.contiguous()


def _threshold(value, name):
  """A scalar threshold or None (vtrace.py:215-220, :228-231)."""
  if value is None:
    return None
  if isinstance(value, torch.Tensor):
    if value.dim() != 0:
      raise ValueError('%s must have rank 0' % name)
    value = value.item()
  elif hasattr(value, 'shape') and tuple(getattr(value, 'shape')) != ():
    raise ValueError('%s must have rank 0' % name)
  return float(value)


def _assert_rank(x, rank, name):
  if x.dim() != rank:
    raise ValueError('%s must have rank %d, got shape %s' %
                     (name, rank, tuple(x.shape)))


def log_probs_from_logits_and_actions(policy_logits, actions):
  """Computes action log-probs from policy logits and actions.

  Args:
    policy_logits: float32 [T, B, NUM_ACTIONS] un-normalized log-probabilities.
    actions: int32 [T, B] actions.

  Returns:
    float32 [T, B]: log-probability of the chosen action under the policy
    (vtrace.py:45-68).  Differentiable w.r.t. `policy_logits`.
  """
  dev = _device_of(policy_logits, actions)
  policy_logits = _to(policy_logits, torch.float32, dev)
  actions = _to(actions, torch.int32, dev)
  _assert_rank(policy_logits, 3, 'policy_logits')
  _assert_rank(actions, 2, 'actions')
  return _ops.LogProbs.apply(policy_logits, actions)


def from_logits(
    behaviour_policy_logits, target_policy_logits, actions,
    discounts, rewards, values, bootstrap_value,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
    name='vtrace_from_logits'):
  """V-trace for softmax policies (vtrace.py:71-161).

  One fused kernel computes the action log-probs of both policies, log_rhos,
  the clipped importance weights and the reverse recurrence in a single pass
  over the [T, B, NUM_ACTIONS] logits.

  Args and returns are those of the reference: see its docstring.  `name` is
  accepted and ignored.  `vs` and `pg_advantages` carry no gradient
  (vtrace.py:279-280); `log_rhos` and the two log-prob fields stay
  differentiable w.r.t. the logits, as in the reference (:156-161).
  """
  del name
  dev = _device_of(behaviour_policy_logits, target_policy_logits, actions,
                   discounts, rewards, values, bootstrap_value)
  behaviour_policy_logits = _to(behaviour_policy_logits, torch.float32, dev)
  target_policy_logits = _to(target_policy_logits, torch.float32, dev)
  actions = _to(actions, torch.int32, dev)
  discounts = _to(discounts, torch.float32, dev)
  rewards = _to(rewards, torch.float32, dev)
  values = _to(values, torch.float32, dev)
  bootstrap_value = _to(bootstrap_value, torch.float32, dev)
  clip_rho_threshold = _threshold(clip_rho_threshold, 'clip_rho_threshold')
  clip_pg_rho_threshold = _threshold(clip_pg_rho_threshold,
                                     'clip_pg_rho_threshold')

  # Make sure tensor ranks are as expected (vtrace.py:134-138, :223-227).
  _assert_rank(behaviour_policy_logits, 3, 'behaviour_policy_logits')
  _assert_rank(target_policy_logits, 3, 'target_policy_logits')
  _assert_rank(actions, 2, 'actions')
  _assert_rank(values, 2, 'values')
  _assert_rank(bootstrap_value, 1, 'bootstrap_value')
  _assert_rank(discounts, 2, 'discounts')
  _assert_rank(rewards, 2, 'rewards')
  T, B, A = target_policy_logits.shape
  if tuple(behaviour_policy_logits.shape) != (T, B, A):
    raise ValueError('behaviour and target logits differ in shape')
  for t, n in ((actions, 'actions'), (discounts, 'discounts'),
               (rewards, 'rewards'), (values, 'values')):
    if tuple(t.shape) != (T, B):
      raise ValueError('%s must have shape [T, B] = %s' % (n, (T, B)))
  if tuple(bootstrap_value.shape) != (B,):
    raise ValueError('bootstrap_value must have shape [B]')

  vs, pg, log_rhos, b_lp, t_lp = _ops.FromLogits.apply(
      behaviour_policy_logits, target_policy_logits, actions, discounts,
      rewards, values, bootstrap_value, clip_rho_threshold,
      clip_pg_rho_threshold, _lib.PATHS['auto'])
  return VTraceFromLogitsReturns(
      vs=vs, pg_advantages=pg, log_rhos=log_rhos,
      behaviour_action_log_probs=b_lp, target_action_log_probs=t_lp)


def from_importance_weights(
    log_rhos, discounts, rewards, values, bootstrap_value,
    clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0,
    name='vtrace_from_importance_weights'):
  """V-trace from log importance weights (vtrace.py:164-280).

  Supports the reference's extra trailing dimensions: all tensors may carry
  the same number of additional dimensions (e.g. rewards [T, B, C], values
  [T, B, C], bootstrap_value [B, C]) and those dimensions broadcast
  (vtrace.py:178-180, vtrace_test.py:249-259).  Returns a `VTraceReturns`
  whose fields carry no gradient (vtrace.py:279-280).
  """
  del name
  # The plain case - five contiguous float32 CUDA tensors, [T, B] and [B] -
  # goes straight through the C++ host layer: the kernel takes about 20 us at
  # B = 32768 and the coercions below twice that.
  plain = (log_rhos, discounts, rewards, values, bootstrap_value)
  if all(isinstance(t, torch.Tensor) and t.is_cuda and t.dtype is torch.float32 and t.is_contiguous()
         for t in plain) and [t.dim() for t in plain] == [2, 2, 2, 2, 1] \
      and not torch.compiler.is_compiling():
    from scalable_agent_b200 import _ext
    ext = _ext.get()
    if ext is not None:
      vs, pg = ext.importance_weights(
          log_rhos.detach(), discounts.detach(), rewards.detach(), values.detach(), bootstrap_value.detach(),
          _threshold(clip_rho_threshold, 'clip_rho_threshold'),
          _threshold(clip_pg_rho_threshold, 'clip_pg_rho_threshold'))
      return VTraceReturns(vs=vs, pg_advantages=pg)
  dev = _device_of(log_rhos, discounts, rewards, values, bootstrap_value)
  log_rhos = _to(log_rhos, torch.float32, dev)
  discounts = _to(discounts, torch.float32, dev)
  rewards = _to(rewards, torch.float32, dev)
  values = _to(values, torch.float32, dev)
  bootstrap_value = _to(bootstrap_value, torch.float32, dev)
  clip_rho_threshold = _threshold(clip_rho_threshold, 'clip_rho_threshold')
  clip_pg_rho_threshold = _threshold(clip_pg_rho_threshold,
                                     'clip_pg_rho_threshold')

  # Make sure tensor ranks are consistent (vtrace.py:222-231).
  rho_rank = log_rhos.dim()
  if rho_rank == 1:
    # One sequence without a batch dimension ([T] tensors, scalar bootstrap):
    # nothing in vtrace.py:222-231 forbids it.  Run it as a batch of one.
    _assert_rank(values, 1, 'values')
    _assert_rank(bootstrap_value, 0, 'bootstrap_value')
    _assert_rank(discounts, 1, 'discounts')
    _assert_rank(rewards, 1, 'rewards')
    out = from_importance_weights(
        log_rhos[:, None], discounts[:, None], rewards[:, None], values[:, None],
        bootstrap_value[None], clip_rho_threshold, clip_pg_rho_threshold)
    return VTraceReturns(vs=out.vs[:, 0], pg_advantages=out.pg_advantages[:, 0])
  if rho_rank < 1:
    raise ValueError('log_rhos must have rank >= 1')
  _assert_rank(values, rho_rank, 'values')
  _assert_rank(bootstrap_value, rho_rank - 1, 'bootstrap_value')
  _assert_rank(discounts, rho_rank, 'discounts')
  _assert_rank(rewards, rho_rank, 'rewards')

  T, B = values.shape[0], values.shape[1]
  for t, n in ((log_rhos, 'log_rhos'), (discounts, 'discounts'),
               (rewards, 'rewards')):
    if tuple(t.shape[:2]) != (T, B):
      raise ValueError('%s must start with [T, B] = %s' % (n, (T, B)))
  if bootstrap_value.shape[0] != B:
    raise ValueError('bootstrap_value must start with [B]')
  trailing = torch.broadcast_shapes(
      log_rhos.shape[2:], discounts.shape[2:], rewards.shape[2:],
      values.shape[2:], bootstrap_value.shape[1:])
  C = int(math.prod(trailing))
  full = (T, B) + tuple(trailing)

  def widen(x, lead):
    shape = lead + tuple(trailing)
    return x if tuple(x.shape) == shape else x.expand(shape).contiguous()

  bcast_mask = 0
  if C > 1 and log_rhos.numel() == T * B:
    bcast_mask |= 1          # trailing dims all 1: index by batch in-kernel
  else:
    log_rhos = widen(log_rhos, (T, B))
  if C > 1 and discounts.numel() == T * B:
    bcast_mask |= 2
  else:
    discounts = widen(discounts, (T, B))
  rewards = widen(rewards, (T, B))
  values = widen(values, (T, B))
  bootstrap_value = widen(bootstrap_value, (B,))

  with torch.no_grad():
    vs, pg = _ops.call(
        _ops.from_importance_weights,
        log_rhos.detach(), discounts.detach(), rewards.detach(),
        values.detach().reshape(T, B, C), bootstrap_value.detach(), C,
        bcast_mask, clip_rho_threshold, clip_pg_rho_threshold)
  return VTraceReturns(vs=vs.reshape(full), pg_advantages=pg.reshape(full))
