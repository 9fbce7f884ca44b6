"""CPU oracle for the optimiser step that follows the hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/`` and the benchmark scripts' CPU-baseline legs may import this
module; the shipped path is the CUDA library under ``scalable_agent_b200/csrc``.

What it restates (float64 numpy):

* ``experiment.py:410-412``  ``tf.train.polynomial_decay(FLAGS.learning_rate,
  num_env_frames, FLAGS.total_environment_frames, 0)`` - linear decay to zero
  with the step clipped at ``decay_steps`` (power = 1, cycle = False);
* ``experiment.py:413-415``  ``tf.train.RMSPropOptimizer(learning_rate,
  FLAGS.decay, FLAGS.momentum, FLAGS.epsilon).minimize(total_loss)`` - the
  non-centred update, slots ``rms`` initialised to ones and ``momentum`` to
  zeros;
* the data-parallel form of it: with the batch sharded over N ranks the loss
  is a sum over columns (``experiment.py:328,335,343``), so the gradient the
  optimiser sees is the SUM of the per-rank gradients.

The arithmetic lives in a third-party dependency that is absent from the
reference tree: TensorFlow, pinned ``tensorflow==1.9.0`` (``Dockerfile:38``).
Its published definition of the op (``ApplyRMSProp``, training_ops) is

    ms  <- rho * ms + (1 - rho) * grad^2
    mom <- momentum * mom + lr * grad / sqrt(ms + epsilon)
    var <- var - mom

and of the schedule (``tf.train.polynomial_decay``)

    step <- min(global_step, decay_steps)
    lr   <- (lr0 - end) * (1 - step / decay_steps)^power + end

**Parity unpinned**: the reference holds no test, golden vector or fixture for
the optimiser, and TensorFlow 1.x cannot be installed here (no wheel for this
Python, no network), so this restatement is the oracle for the row.
"""
import numpy as np


def polynomial_decay(learning_rate, global_step, decay_steps,
                     end_learning_rate=0.0, power=1.0):
  """experiment.py:411-412 (tf.train.polynomial_decay, cycle=False)."""
  step = min(float(global_step), float(decay_steps))
  frac = 1.0 - step / float(decay_steps)
  return (float(learning_rate) - float(end_learning_rate)) * frac ** float(power) \
      + float(end_learning_rate)


def rmsprop_slots(var):
  """Initial slots of tf.train.RMSPropOptimizer: rms = 1, momentum = 0."""
  var = np.asarray(var, dtype=np.float64)
  return np.ones_like(var), np.zeros_like(var)


def rmsprop_step(var, ms, mom, grad, learning_rate, decay, momentum, epsilon):
  """One ApplyRMSProp (experiment.py:413-415).  Returns new (var, ms, mom)."""
  var, ms, mom, grad = (np.asarray(x, dtype=np.float64) for x in (var, ms, mom, grad))
  ms = decay * ms + (1.0 - decay) * grad * grad
  mom = momentum * mom + learning_rate * grad / np.sqrt(ms + epsilon)
  return var - mom, ms, mom


def data_parallel_step(var, ms, mom, rank_grads, learning_rate, decay, momentum,
                       epsilon):
  """The same step when the batch is sharded: the gradient is the sum over ranks."""
  total = np.zeros_like(np.asarray(var, dtype=np.float64))
  for g in rank_grads:              # rank order, as the kernel sums
    total = total + np.asarray(g, dtype=np.float64)
  return rmsprop_step(var, ms, mom, total, learning_rate, decay, momentum, epsilon)


def slice_bounds(n, world, rank):
  """[lo, hi) of the parameters rank `rank` updates: equal runs of whole
  4-element groups (the kernel moves float4s), the last ranks possibly short."""
  groups = (n + 3) // 4
  per = ((groups + world - 1) // world) * 4
  lo = min(n, rank * per)
  return lo, min(n, lo + per)
