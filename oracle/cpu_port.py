"""Op-for-op float32 CPU port of the reference path.  TEST / BASELINE INFRASTRUCTURE ONLY.

The reference pins V-trace to the host (`experiment.py:389`,
``with tf.device('/cpu')``) and lets TF1 autodiff produce the backward
(`experiment.py:415`).  TensorFlow 1.x cannot be installed in this image, so
this file restates that CPU path with torch-CPU float32 ops, one torch op per
TF op and a Python loop of T iterations where the reference has
``tf.scan(parallel_iterations=1)`` (`vtrace.py:247-262`), with autograd for
the backward.  It uses every host thread torch is given.  It is what
``bench.py --impl reference`` and ``bench.py``'s ``cpu_baseline`` time; it is
never on the product path.
"""
import torch
import torch.nn.functional as F


def _log_probs(policy_logits, actions):
  """vtrace.py:45-68."""
  T, B, A = policy_logits.shape
  ce = F.cross_entropy(policy_logits.reshape(T * B, A),
                       actions.reshape(T * B).long(), reduction='none')
  return -ce.reshape(T, B)


def from_importance_weights(log_rhos, discounts, rewards, values,
                            bootstrap_value, clip_rho_threshold=1.0,
                            clip_pg_rho_threshold=1.0):
  """vtrace.py:164-280."""
  rhos = torch.exp(log_rhos)
  if clip_rho_threshold is not None:
    clipped_rhos = torch.clamp(rhos, max=float(clip_rho_threshold))
  else:
    clipped_rhos = rhos
  cs = torch.clamp(rhos, max=1.0)
  values_t_plus_1 = torch.cat([values[1:], bootstrap_value[None]], 0)
  deltas = clipped_rhos * (rewards + discounts * values_t_plus_1 - values)
  with torch.no_grad():                       # back_prop=False, vtrace.py:260
    acc = torch.zeros_like(bootstrap_value)
    result = []
    for t in range(discounts.shape[0] - 1, -1, -1):
      acc = deltas[t] + discounts[t] * cs[t] * acc
      result.append(acc)
    result.reverse()
    vs_minus_v_xs = torch.stack(result)
  vs = vs_minus_v_xs + values
  vs_t_plus_1 = torch.cat([vs[1:], bootstrap_value[None]], 0)
  if clip_pg_rho_threshold is not None:
    clipped_pg_rhos = torch.clamp(rhos, max=float(clip_pg_rho_threshold))
  else:
    clipped_pg_rhos = rhos
  pg_advantages = clipped_pg_rhos * (
      rewards + discounts * vs_t_plus_1 - values)
  return vs.detach(), pg_advantages.detach()


def learner_loss(behaviour_policy_logits, target_policy_logits, actions,
                 discounts, rewards, values, bootstrap_value,
                 baseline_cost=0.5, entropy_cost=0.00025,
                 clip_rho_threshold=1.0, clip_pg_rho_threshold=1.0):
  """experiment.py:389-407 forward, then backward into logits and values.

  Returns (losses[3], d_target_logits, d_values, vs, pg_advantages).
  """
  target = target_policy_logits.detach().requires_grad_(True)
  baseline = values.detach().requires_grad_(True)
  target_lp = _log_probs(target, actions)
  behaviour_lp = _log_probs(behaviour_policy_logits, actions)
  log_rhos = target_lp - behaviour_lp
  vs, pg_adv = from_importance_weights(
      log_rhos, discounts, rewards, baseline, bootstrap_value,
      clip_rho_threshold, clip_pg_rho_threshold)
  # compute_policy_gradient_loss, experiment.py:338-343
  T, B, A = target.shape
  ce = F.cross_entropy(target.reshape(T * B, A), actions.reshape(T * B).long(),
                       reduction='none').reshape(T, B)
  pg_loss = torch.sum(ce * pg_adv)
  # compute_baseline_loss, experiment.py:324-328
  baseline_loss = 0.5 * torch.sum(torch.square(vs - baseline))
  # compute_entropy_loss, experiment.py:331-335
  policy = F.softmax(target, dim=-1)
  log_policy = F.log_softmax(target, dim=-1)
  entropy_loss = -torch.sum(torch.sum(-policy * log_policy, dim=-1))
  total = pg_loss + baseline_cost * baseline_loss + entropy_cost * entropy_loss
  total.backward()
  losses = torch.stack([pg_loss.detach(), baseline_loss.detach(),
                        entropy_loss.detach()])
  return losses, target.grad, baseline.grad, vs, pg_adv
