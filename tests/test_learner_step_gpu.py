"""One whole learner step through the library, as build_learner strings it
together (experiment.py:362-415): heads on the core output (:200-203) ->
V-trace + the three losses on [T+1, ...] tensors (:366-407) -> backward ->
polynomial-decay RMSProp (:410-415).

The heads and their backward are torch (a plain library contraction, outside
the path); the loss block and the optimiser are this library's kernels.  The
oracle redoes the same step in float64 numpy: logits / baseline from the same
weights, oracle.fused_vtrace_loss for the loss and its gradients w.r.t. logits
and baseline, the chain rule by hand into the head weights, and
rmsprop_oracle for the update.
"""
import numpy as np
import pytest
import torch

from oracle import rmsprop_oracle
from oracle import vtrace_oracle as oracle
from tests.conftest import assert_close

pytestmark = pytest.mark.gpu

T, B, A, H = 20, 48, 9, 32
BC, EC, GAMMA = 0.5, 0.00025, 0.99               # experiment.py:70-72
LR, DECAY, MOMENTUM, EPSILON = 0.00048, 0.99, 0.0, 0.1   # experiment.py:92-95
FRAMES_PER_STEP = B * T * 4                      # experiment.py:420-421


def test_learner_step_matches_float64_oracle():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import experiment, optim
  rng = np.random.default_rng(11)
  core = rng.standard_normal((T + 1, B, H)).astype(np.float32)
  behaviour = rng.standard_normal((T + 1, B, A)).astype(np.float32)
  actions = rng.integers(0, A, (T + 1, B)).astype(np.int32)
  rewards = (2 * rng.standard_normal((T + 1, B))).astype(np.float32)
  done = rng.random((T + 1, B)) < 0.05
  w_pi = (rng.standard_normal((H, A)) / np.sqrt(H)).astype(np.float32)
  b_pi = np.zeros(A, np.float32)
  w_v = (rng.standard_normal((H, 1)) / np.sqrt(H)).astype(np.float32)
  b_v = np.zeros(1, np.float32)

  dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
  params = [torch.nn.Parameter(dev(p)) for p in (w_pi, b_pi, w_v, b_v)]
  frames = [0]
  opt = optim.RMSPropOptimizer(
      params, lambda: optim.polynomial_decay(LR, frames[0], 10 * FRAMES_PER_STEP, 0),
      DECAY, MOMENTUM, EPSILON)
  core_t, behaviour_t = dev(core), dev(behaviour)
  actions_t, rewards_t, done_t = dev(actions), dev(rewards), dev(done)

  shadow = [p.astype(np.float64) for p in (w_pi, b_pi, w_v, b_v)]
  slots = [rmsprop_oracle.rmsprop_slots(s) for s in shadow]
  core64 = core.astype(np.float64)
  for step in range(3):
    # --- the library ---
    opt.zero_grad()
    logits = core_t @ params[0] + params[1]                       # :200-201
    baseline = (core_t @ params[2] + params[3]).squeeze(-1)       # :202-203
    out = experiment.compute_learner_loss(
        behaviour_t, actions_t, rewards_t, done_t, logits, baseline,
        discounting=GAMMA, reward_clipping='abs_one', baseline_cost=BC, entropy_cost=EC)
    out.total_loss.backward()
    opt.step()

    # --- the oracle, from the parameters the step started with ---
    o_logits = core64 @ shadow[0] + shadow[1]
    o_baseline = (core64 @ shadow[2] + shadow[3])[..., 0]
    ref = oracle.fused_vtrace_loss(
        behaviour[1:], o_logits[:-1], actions[1:], None, rewards[1:], o_baseline[:-1],
        o_baseline[-1], 1.0, 1.0, BC, EC, reward_clipping='abs_one', done=done[1:],
        discounting=GAMMA)
    assert_close(out.total_loss.detach().cpu().numpy(), ref.total_loss, 1e-5, 'total loss')
    x = core64[:-1].reshape(-1, H)                 # the last step only feeds the bootstrap
    d_logits = ref.d_target_logits.reshape(-1, A)
    d_values = ref.d_values.reshape(-1, 1)
    grads = [x.T @ d_logits, d_logits.sum(0), x.T @ d_values, d_values.sum(0)]
    lr = rmsprop_oracle.polynomial_decay(LR, frames[0], 10 * FRAMES_PER_STEP, 0)
    for i, g in enumerate(grads):
      v, ms, mom = rmsprop_oracle.rmsprop_step(shadow[i], slots[i][0], slots[i][1], g, lr, DECAY,
                                               MOMENTUM, EPSILON)
      shadow[i], slots[i] = v, (ms, mom)
    frames[0] += FRAMES_PER_STEP
    for i, p in enumerate(params):
      assert_close(p.detach().cpu().numpy(), shadow[i], 1e-5, 'step %d parameter %d' % (step, i))
      # follow the device trajectory so float32 rounding of the heads does not pile up
      shadow[i] = p.detach().cpu().numpy().astype(np.float64)
  # the update moved the weights by about lr / sqrt(ms + eps) per step: not a no-op
  assert float(np.abs(shadow[0] - w_pi).max()) > 1e-4
