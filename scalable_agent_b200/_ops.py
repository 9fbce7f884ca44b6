"""PyTorch custom ops over the C ABI (include/vtrace_b200.h).

PyTorch is plumbing here: it owns device memory and streams and hands raw
pointers to libvtrace_b200.so.  Every op below launches hand-written sm_100a
kernels; none has a CPU or eager-PyTorch implementation, and calling one
without a CUDA device raises.
"""
import ctypes

import torch

from scalable_agent_b200 import _lib

_NS = 'b200vtrace'


_cuda_ok = False


def require_cuda():
  global _cuda_ok
  if _cuda_ok:              # asked once: the answer does not change while we run
    return
  if not torch.cuda.is_available():
    raise RuntimeError(
        'scalable_agent_b200 needs a CUDA device (B200, sm_100a); '
        'there is no CPU fallback.')
  _lib.load()
  _cuda_ok = True


def _stream(device):
  """The current stream of `device` as the void* the C ABI takes."""
  raw = getattr(torch._C, '_cuda_getCurrentRawStream', None)
  if raw is not None:       # same stream, without building a torch.cuda.Stream
    return ctypes.c_void_p(raw(device.index if device.index is not None
                               else torch.cuda.current_device()))
  return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


# Every kernel entry point below is registered as a torch custom op (so that
# torch.compile / export see an opaque, shape-annotated op), but a call through
# the dispatcher costs 50-100 us of host time - more than most of the kernels
# take.  The autograd wrappers and the API therefore call the op's own Python
# body directly in eager mode and the registered op only while compiling.
_BODY = {}


def _custom_op(name, mutates_args=()):
  def register(fn):
    op = torch.library.custom_op(_NS + '::' + name, mutates_args=mutates_args)(fn)
    _BODY[op] = fn
    return op
  return register


def call(op, *args):
  """`op(*args)`; in eager mode without the trip through the dispatcher."""
  if torch.compiler.is_compiling():
    return op(*args)
  return _BODY[op](*args)


class _NoSwitch(object):
  def __enter__(self):
    return None

  def __exit__(self, *exc):
    return False


_NO_SWITCH = _NoSwitch()


def _on(device):
  """Context that makes `device` current for the launch; nothing to do - and
  nothing spent - when it already is."""
  if device.index is None or device.index == torch.cuda.current_device():
    return _NO_SWITCH
  return torch.cuda.device(device)


def _ptr(t):
  return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _check_cuda(*tensors):
  dev = None
  for t in tensors:
    if t is None:
      continue
    if not t.is_cuda:
      raise RuntimeError('b200vtrace ops take CUDA tensors only')
    if dev is None:
      dev = t.device
    elif t.device != dev:
      raise RuntimeError('b200vtrace ops need all tensors on one device')
    if not t.is_contiguous():
      raise RuntimeError('b200vtrace ops need contiguous tensors')
  return dev


# ---------------------------------------------------------------------------
# Fused V-trace + losses + gradients (vtr_from_logits_fused_f32)
# ---------------------------------------------------------------------------
@_custom_op('from_logits_fused')
def from_logits_fused(
    behaviour_policy_logits: torch.Tensor, target_policy_logits: torch.Tensor,
    actions: torch.Tensor, discounts: torch.Tensor, rewards: torch.Tensor,
    values: torch.Tensor, bootstrap_value: torch.Tensor,
    done: torch.Tensor | None,
    clip_rho_threshold: float | None, clip_pg_rho_threshold: float | None,
    baseline_cost: float, entropy_cost: float, discounting: float,
    reward_clip_mode: int, want_aux: bool, want_losses: bool,
    want_grads: bool, path: int, pad_grads: bool = False
) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor,
           torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
  """Returns (vs, pg_adv, log_rhos, b_lp, t_lp, losses[5], d_logits, d_values);
  losses = [pg, baseline, entropy, total, total]: the last element is the
  weighted total once more, written by the kernel into a slot of its own
  (vtr_fused_params.total_loss) so that a caller can hand out a
  differentiable total that does not alias the constant components.

  Outputs that were not asked for come back as empty tensors.  With
  `pad_grads` the two gradient tensors get one extra, zeroed time step at the
  end ([T+1, ...]): the gradient w.r.t. a learner output of T+1 steps whose
  last step only feeds the bootstrap value (experiment.py:366-375).
  """
  require_cuda()
  dev = _check_cuda(behaviour_policy_logits, target_policy_logits, actions,
                    discounts, rewards, values, bootstrap_value, done)
  T, B, A = target_policy_logits.shape
  f32 = dict(dtype=torch.float32, device=dev)
  empty = lambda: torch.empty(0, **f32)
  vs = torch.empty((T, B), **f32)
  pg = torch.empty((T, B), **f32)
  log_rhos = torch.empty((T, B), **f32) if want_aux else empty()
  b_lp = torch.empty((T, B), **f32) if want_aux else empty()
  t_lp = torch.empty((T, B), **f32) if want_aux else empty()
  losses = torch.empty(5, **f32) if want_losses else empty()
  Tg = T + 1 if pad_grads else T
  d_logits = torch.empty((Tg, B, A), **f32) if want_grads else empty()
  d_values = torch.empty((Tg, B), **f32) if want_grads else empty()
  if want_grads and pad_grads:
    d_logits[T].zero_()
    d_values[T].zero_()

  lib = _lib.load()
  p = _lib.FusedParams()
  p.T, p.B, p.A = T, B, A
  p.flags = ((_lib.FLAG_NO_CLIP_RHO if clip_rho_threshold is None else 0) |
             (_lib.FLAG_NO_CLIP_PG_RHO if clip_pg_rho_threshold is None else 0))
  p.behaviour_policy_logits = behaviour_policy_logits.data_ptr()
  p.target_policy_logits = target_policy_logits.data_ptr()
  p.actions = actions.data_ptr()
  p.discounts = discounts.data_ptr() if discounts.numel() else None
  p.rewards = rewards.data_ptr()
  p.values = values.data_ptr()
  p.bootstrap_value = bootstrap_value.data_ptr()
  p.done = done.data_ptr() if done is not None else None
  p.clip_rho_threshold = 0.0 if clip_rho_threshold is None else clip_rho_threshold
  p.clip_pg_rho_threshold = (0.0 if clip_pg_rho_threshold is None
                             else clip_pg_rho_threshold)
  p.baseline_cost, p.entropy_cost = baseline_cost, entropy_cost
  p.discounting = discounting
  p.reward_clip_mode = reward_clip_mode
  p.path = path
  p.vs, p.pg_advantages = vs.data_ptr(), pg.data_ptr()
  if want_aux:
    p.log_rhos = log_rhos.data_ptr()
    p.behaviour_action_log_probs = b_lp.data_ptr()
    p.target_action_log_probs = t_lp.data_ptr()
  if want_grads:
    p.d_target_logits, p.d_values = d_logits.data_ptr(), d_values.data_ptr()
  workspace, ws_bytes = None, 0
  if want_losses:
    p.losses = losses.data_ptr()
    p.total_loss = losses.data_ptr() + 16
    ws_bytes = lib.vtr_fused_workspace_bytes(T, B, A)
    workspace = torch.empty(ws_bytes // 8, dtype=torch.float64, device=dev)
  with _on(dev):
    rc = lib.vtr_from_logits_fused_f32(
        ctypes.byref(p), _ptr(workspace), ws_bytes, _stream(dev))
  _lib.check(rc, 'vtr_from_logits_fused_f32')
  return vs, pg, log_rhos, b_lp, t_lp, losses, d_logits, d_values


@from_logits_fused.register_fake
def _(behaviour_policy_logits, target_policy_logits, actions, discounts,
      rewards, values, bootstrap_value, done, clip_rho_threshold,
      clip_pg_rho_threshold, baseline_cost, entropy_cost, discounting,
      reward_clip_mode, want_aux, want_losses, want_grads, path,
      pad_grads=False):
  T, B, A = target_policy_logits.shape
  new = target_policy_logits.new_empty
  tb = lambda on: new((T, B)) if on else new((0,))
  Tg = T + 1 if pad_grads else T
  return (new((T, B)), new((T, B)), tb(want_aux), tb(want_aux), tb(want_aux),
          new((5,)) if want_losses else new((0,)),
          new((Tg, B, A)) if want_grads else new((0,)),
          new((Tg, B)) if want_grads else new((0,)))


# ---------------------------------------------------------------------------
# Stand-alone pieces
# ---------------------------------------------------------------------------
@_custom_op('from_importance_weights')
def from_importance_weights(
    log_rhos: torch.Tensor, discounts: torch.Tensor, rewards: torch.Tensor,
    values: torch.Tensor, bootstrap_value: torch.Tensor, C: int,
    bcast_mask: int, clip_rho_threshold: float | None,
    clip_pg_rho_threshold: float | None
) -> tuple[torch.Tensor, torch.Tensor]:
  """values/rewards are [T,B,C] flattened views; see vtr_from_importance_weights_f32."""
  require_cuda()
  dev = _check_cuda(log_rhos, discounts, rewards, values, bootstrap_value)
  T, B = values.shape[0], values.shape[1]
  vs = torch.empty_like(values)
  pg = torch.empty_like(values)
  flags = ((_lib.FLAG_NO_CLIP_RHO if clip_rho_threshold is None else 0) |
           (_lib.FLAG_NO_CLIP_PG_RHO if clip_pg_rho_threshold is None else 0))
  with _on(dev):
    rc = _lib.load().vtr_from_importance_weights_f32(
        _ptr(log_rhos), _ptr(discounts), _ptr(rewards), _ptr(values),
        _ptr(bootstrap_value), T, B, C, bcast_mask,
        0.0 if clip_rho_threshold is None else clip_rho_threshold,
        0.0 if clip_pg_rho_threshold is None else clip_pg_rho_threshold,
        flags, _ptr(vs), _ptr(pg), _stream(dev))
  _lib.check(rc, 'vtr_from_importance_weights_f32')
  return vs, pg


@from_importance_weights.register_fake
def _(log_rhos, discounts, rewards, values, bootstrap_value, C, bcast_mask,
      clip_rho_threshold, clip_pg_rho_threshold):
  return torch.empty_like(values), torch.empty_like(values)


@_custom_op('log_probs')
def log_probs(policy_logits: torch.Tensor, actions: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(policy_logits, actions)
  A = policy_logits.shape[-1]
  rows = policy_logits.numel() // A
  out = torch.empty(policy_logits.shape[:-1], dtype=torch.float32, device=dev)
  with _on(dev):
    rc = _lib.load().vtr_log_probs_from_logits_and_actions_f32(
        _ptr(policy_logits), _ptr(actions), rows, A, _ptr(out), _stream(dev))
  _lib.check(rc, 'vtr_log_probs_from_logits_and_actions_f32')
  return out


@log_probs.register_fake
def _(policy_logits, actions):
  return policy_logits.new_empty(policy_logits.shape[:-1])


def _loss_workspace(units, dev):
  nbytes = _lib.load().vtr_loss_workspace_bytes(units)
  return torch.empty(nbytes // 8, dtype=torch.float64, device=dev), nbytes


@_custom_op('pg_loss')
def pg_loss(logits: torch.Tensor, actions: torch.Tensor,
            advantages: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(logits, actions, advantages)
  A = logits.shape[-1]
  rows = logits.numel() // A
  loss = torch.empty(1, dtype=torch.float32, device=dev)
  ws, nbytes = _loss_workspace(rows, dev)
  with _on(dev):
    rc = _lib.load().vtr_policy_gradient_loss_f32(
        _ptr(logits), _ptr(actions), _ptr(advantages), rows, A, _ptr(loss),
        _ptr(ws), nbytes, _stream(dev))
  _lib.check(rc, 'vtr_policy_gradient_loss_f32')
  return loss.reshape(())


@pg_loss.register_fake
def _(logits, actions, advantages):
  return logits.new_empty(())


@_custom_op('pg_loss_bwd')
def pg_loss_bwd(logits: torch.Tensor, actions: torch.Tensor,
                advantages: torch.Tensor, grad_out: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(logits, actions, advantages, grad_out)
  A = logits.shape[-1]
  d_logits = torch.empty_like(logits)
  with _on(dev):
    rc = _lib.load().vtr_policy_gradient_loss_bwd_f32(
        _ptr(logits), _ptr(actions), _ptr(advantages), _ptr(grad_out),
        logits.numel() // A, A, _ptr(d_logits), _stream(dev))
  _lib.check(rc, 'vtr_policy_gradient_loss_bwd_f32')
  return d_logits


@pg_loss_bwd.register_fake
def _(logits, actions, advantages, grad_out):
  return torch.empty_like(logits)


@_custom_op('entropy_loss')
def entropy_loss(logits: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(logits)
  A = logits.shape[-1]
  rows = logits.numel() // A
  loss = torch.empty(1, dtype=torch.float32, device=dev)
  ws, nbytes = _loss_workspace(rows, dev)
  with _on(dev):
    rc = _lib.load().vtr_entropy_loss_f32(
        _ptr(logits), rows, A, _ptr(loss), _ptr(ws), nbytes, _stream(dev))
  _lib.check(rc, 'vtr_entropy_loss_f32')
  return loss.reshape(())


@entropy_loss.register_fake
def _(logits):
  return logits.new_empty(())


@_custom_op('entropy_loss_bwd')
def entropy_loss_bwd(logits: torch.Tensor, grad_out: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(logits, grad_out)
  A = logits.shape[-1]
  d_logits = torch.empty_like(logits)
  with _on(dev):
    rc = _lib.load().vtr_entropy_loss_bwd_f32(
        _ptr(logits), _ptr(grad_out), logits.numel() // A, A, _ptr(d_logits),
        _stream(dev))
  _lib.check(rc, 'vtr_entropy_loss_bwd_f32')
  return d_logits


@entropy_loss_bwd.register_fake
def _(logits, grad_out):
  return torch.empty_like(logits)


@_custom_op('baseline_loss')
def baseline_loss(advantages: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(advantages)
  n = advantages.numel()
  loss = torch.empty(1, dtype=torch.float32, device=dev)
  ws, nbytes = _loss_workspace(n, dev)
  with _on(dev):
    rc = _lib.load().vtr_baseline_loss_f32(
        _ptr(advantages), n, _ptr(loss), _ptr(ws), nbytes, _stream(dev))
  _lib.check(rc, 'vtr_baseline_loss_f32')
  return loss.reshape(())


@baseline_loss.register_fake
def _(advantages):
  return advantages.new_empty(())


@_custom_op('baseline_loss_bwd')
def baseline_loss_bwd(advantages: torch.Tensor,
                      grad_out: torch.Tensor) -> torch.Tensor:
  require_cuda()
  dev = _check_cuda(advantages, grad_out)
  d_adv = torch.empty_like(advantages)
  with _on(dev):
    rc = _lib.load().vtr_baseline_loss_bwd_f32(
        _ptr(advantages), _ptr(grad_out), advantages.numel(), _ptr(d_adv),
        _stream(dev))
  _lib.check(rc, 'vtr_baseline_loss_bwd_f32')
  return d_adv


@baseline_loss_bwd.register_fake
def _(advantages, grad_out):
  return torch.empty_like(advantages)


@_custom_op('scale_inplace', mutates_args=('x',))
def scale_inplace(x: torch.Tensor, scale: torch.Tensor) -> None:
  """x *= scale (device scalar), a no-op launch when scale == 1."""
  require_cuda()
  dev = _check_cuda(x, scale)
  with _on(dev):
    rc = _lib.load().vtr_scale_inplace_f32(
        _ptr(x), _ptr(scale), x.numel(), _stream(dev))
  _lib.check(rc, 'vtr_scale_inplace_f32')


def _scalar_f32(g):
  """Upstream gradient of a scalar loss as a contiguous device float[1]."""
  return g.detach().to(torch.float32).reshape(1).contiguous()


# ---------------------------------------------------------------------------
# Autograd wrappers
# ---------------------------------------------------------------------------
class LogProbs(torch.autograd.Function):
  """log pi(a) with d/dlogits = g * (onehot(a) - softmax)."""

  @staticmethod
  def forward(ctx, policy_logits, actions):
    ctx.save_for_backward(policy_logits, actions)
    return call(log_probs, policy_logits, actions)

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g):
    policy_logits, actions = ctx.saved_tensors
    one = torch.ones(1, dtype=torch.float32, device=g.device)
    # d(sum g*lp) = -(d sum CE*g): reuse the policy-gradient backward kernel.
    return call(pg_loss_bwd, policy_logits, actions, (-g).contiguous(), one), None


class PolicyGradientLoss(torch.autograd.Function):

  @staticmethod
  def forward(ctx, logits, actions, advantages):
    ctx.save_for_backward(logits, actions, advantages)
    return call(pg_loss, logits, actions, advantages)

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g):
    logits, actions, advantages = ctx.saved_tensors
    # advantages are stop_gradient in the reference (experiment.py:341).
    return call(pg_loss_bwd, logits, actions, advantages, _scalar_f32(g)), None, None


class EntropyLoss(torch.autograd.Function):

  @staticmethod
  def forward(ctx, logits):
    ctx.save_for_backward(logits)
    return call(entropy_loss, logits)

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g):
    (logits,) = ctx.saved_tensors
    return call(entropy_loss_bwd, logits, _scalar_f32(g))


class BaselineLoss(torch.autograd.Function):

  @staticmethod
  def forward(ctx, advantages):
    ctx.save_for_backward(advantages)
    return call(baseline_loss, advantages)

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g):
    (advantages,) = ctx.saved_tensors
    return call(baseline_loss_bwd, advantages, _scalar_f32(g))


class FromLogits(torch.autograd.Function):
  """vtrace.from_logits: vs / pg_advantages are constants (vtrace.py:279-280);
  log_rhos and the two action log-prob outputs stay differentiable
  (vtrace.py:156-161)."""

  @staticmethod
  def forward(ctx, behaviour_policy_logits, target_policy_logits, actions,
              discounts, rewards, values, bootstrap_value, clip_rho_threshold,
              clip_pg_rho_threshold, path):
    vs, pg, log_rhos, b_lp, t_lp, _, _, _ = call(
        from_logits_fused,
        behaviour_policy_logits, target_policy_logits, actions, discounts,
        rewards, values, bootstrap_value, None, clip_rho_threshold,
        clip_pg_rho_threshold, 0.0, 0.0, 0.0, 0, True, False, False, path)
    ctx.save_for_backward(behaviour_policy_logits, target_policy_logits, actions)
    ctx.mark_non_differentiable(vs, pg)
    # outputs nobody differentiated through come to backward() as None, not
    # as freshly zero-filled [T, B] tensors
    ctx.set_materialize_grads(False)
    return vs, pg, log_rhos, b_lp, t_lp

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g_vs, g_pg, g_log_rhos, g_b_lp, g_t_lp):
    behaviour, target, actions = ctx.saved_tensors
    one = torch.ones(1, dtype=torch.float32, device=target.device)

    def combine(a, sa, b, sb):
      """sa*a + sb*b over gradients that may be None (= zero)."""
      parts = [s * g if s != 1 else g for g, s in ((a, sa), (b, sb)) if g is not None]
      if not parts:
        return None
      return (parts[0] if len(parts) == 1 else parts[0] + parts[1]).contiguous()

    d_behaviour = d_target = None
    if ctx.needs_input_grad[0]:
      w = combine(g_log_rhos, 1, g_b_lp, -1)
      if w is not None:
        d_behaviour = call(pg_loss_bwd, behaviour, actions, w, one)
    if ctx.needs_input_grad[1]:
      w = combine(g_log_rhos, -1, g_t_lp, -1)
      if w is not None:
        d_target = call(pg_loss_bwd, target, actions, w, one)
    return (d_behaviour, d_target) + (None,) * 8


class FusedLoss(torch.autograd.Function):
  """The whole learner loss of experiment.py:389-407 in one pass.

  The forward launch already produced d(total)/d(target_logits) and
  d(total)/d(values); backward hands them out, scaled in place by the upstream
  gradient only when that is not exactly 1.
  """

  @staticmethod
  def forward(ctx, behaviour_policy_logits, target_policy_logits, actions,
              discounts, rewards, values, bootstrap_value, done,
              clip_rho_threshold, clip_pg_rho_threshold, baseline_cost,
              entropy_cost, discounting, reward_clip_mode, want_aux, path):
    want_grads = ctx.needs_input_grad[1] or ctx.needs_input_grad[5]
    vs, pg, log_rhos, b_lp, t_lp, losses, d_logits, d_values = call(
        from_logits_fused,
        behaviour_policy_logits, target_policy_logits, actions, discounts,
        rewards, values, bootstrap_value, done, clip_rho_threshold,
        clip_pg_rho_threshold, baseline_cost, entropy_cost, discounting,
        reward_clip_mode, want_aux, True, want_grads, path)
    ctx.save_for_backward(d_logits, d_values)
    total = losses[4].clone()
    ctx.backward_ran = False
    ctx.set_materialize_grads(False)
    ctx.mark_non_differentiable(vs, pg, log_rhos, b_lp, t_lp, losses)
    return total, losses, vs, pg, log_rhos, b_lp, t_lp

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g_total, *unused):
    if ctx.backward_ran:
      raise RuntimeError('FusedLoss hands out its stashed gradients in place; '
                         'backward can run only once per forward')
    ctx.backward_ran = True
    d_logits, d_values = ctx.saved_tensors
    if g_total is None:
      return (None,) * 16
    if d_logits.numel():       # empty: no differentiated input asked for a gradient
      g = _scalar_f32(g_total)
      call(scale_inplace, d_logits, g)
      call(scale_inplace, d_values, g)
    grads = [None] * 16
    if ctx.needs_input_grad[1]:
      grads[1] = d_logits
    if ctx.needs_input_grad[5]:
      grads[5] = d_values
    return tuple(grads)


class FusedLearnerLoss(torch.autograd.Function):
  """experiment.py:362-407 on the learner's own [T+1, ...] tensors.

  Takes the learner outputs whole (policy_logits [T+1,B,A], baseline [T+1,B]);
  the shifts of experiment.py:366-375 are views, and the gradients come back
  already shaped [T+1, ...] with a zero last step, so autograd never has to
  pad a [T, ...] gradient into a fresh [T+1, ...] tensor.
  """

  @staticmethod
  def forward(ctx, learner_logits, learner_baseline, behaviour_logits, actions,
              rewards, done, clip_rho_threshold, clip_pg_rho_threshold,
              baseline_cost, entropy_cost, discounting, reward_clip_mode,
              want_aux, path):
    T = learner_logits.shape[0] - 1
    want_grads = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
    empty = torch.empty(0, dtype=torch.float32, device=learner_logits.device)
    vs, pg, log_rhos, b_lp, t_lp, losses, d_logits, d_values = call(
        from_logits_fused,
        behaviour_logits[1:], learner_logits[:T], actions[1:], empty,
        rewards[1:], learner_baseline[:T], learner_baseline[T], done[1:],
        clip_rho_threshold, clip_pg_rho_threshold, baseline_cost, entropy_cost,
        discounting, reward_clip_mode, want_aux, True, want_grads, path, True)
    ctx.save_for_backward(d_logits, d_values)
    ctx.backward_ran = False
    total = losses[4].clone()
    ctx.set_materialize_grads(False)
    ctx.mark_non_differentiable(vs, pg, log_rhos, b_lp, t_lp, losses)
    return total, losses, vs, pg, log_rhos, b_lp, t_lp

  @staticmethod
  @torch.autograd.function.once_differentiable
  def backward(ctx, g_total, *unused):
    if ctx.backward_ran:
      raise RuntimeError('FusedLearnerLoss hands out its stashed gradients in '
                         'place; backward can run only once per forward')
    ctx.backward_ran = True
    d_logits, d_values = ctx.saved_tensors
    if g_total is None:
      return (None,) * 14
    if d_logits.numel():
      g = _scalar_f32(g_total)
      call(scale_inplace, d_logits, g)
      call(scale_inplace, d_values, g)
    grads = [None] * 14
    if ctx.needs_input_grad[0]:
      grads[0] = d_logits
    if ctx.needs_input_grad[1]:
      grads[1] = d_values
    return tuple(grads)
