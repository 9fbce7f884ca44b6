"""The optimiser step that follows the path: `/root/reference/experiment.py:410-415`.

    learning_rate = tf.train.polynomial_decay(FLAGS.learning_rate, num_env_frames,
                                              FLAGS.total_environment_frames, 0)
    optimizer = tf.train.RMSPropOptimizer(learning_rate, FLAGS.decay,
                                          FLAGS.momentum, FLAGS.epsilon)
    train_op = optimizer.minimize(total_loss)

`polynomial_decay` and `RMSPropOptimizer` keep the names, argument order and
defaults of the TensorFlow 1.x functions the reference calls; the optimiser
works on torch CUDA parameters whose gradients autograd has filled in.  The
arithmetic is TensorFlow's ApplyRMSProp (slots `rms` = 1 and `momentum` = 0 at
the start, epsilon inside the square root), done by this library's kernels
(`csrc/vtrace_optim.cu`); there is no PyTorch or CPU fallback.

With `group=` (a `torch.distributed` group over the GPUs of one box) the
optimiser is the data-parallel learner's: every rank back-propagates its shard
of the batch, and one kernel per rank sums the gradients over NVLink peer
memory, updates that rank's slice of the parameters and stores it into every
rank's copy (`vtr_rmsprop_allreduce_f32`).
"""
import ctypes

import torch

from scalable_agent_b200 import _lib


def polynomial_decay(learning_rate, global_step, decay_steps,
                     end_learning_rate=0.0001, power=1.0, cycle=False, name=None):
  """tf.train.polynomial_decay (same arguments and defaults).

  The reference passes `end_learning_rate=0` positionally
  (`experiment.py:411-412`).  `cycle=True` is not used by it and not offered.
  """
  del name
  if cycle:
    raise ValueError('polynomial_decay: cycle=True is not supported')
  return _lib.load().vtr_polynomial_decay(
      float(learning_rate), int(global_step), int(decay_steps),
      float(end_learning_rate), float(power))


def _flatten(params):
  """Offsets of the parameters in one flat buffer, each starting on a
  16-byte boundary so that the views stay usable by vectorised kernels."""
  offsets, total = [], 0
  for p in params:
    offsets.append(total)
    total += (p.numel() + 3) // 4 * 4
  return offsets, total


class RMSPropOptimizer(object):
  """tf.train.RMSPropOptimizer(learning_rate, decay, momentum, epsilon) for torch.

  Args:
    params: iterable of float32 CUDA parameters (all on one device).  They are
      moved into one flat buffer and become views of it; so do their `.grad`s,
      which autograd then accumulates into in place.
    learning_rate: float, or a callable returning one (evaluated every step),
      e.g. `lambda: polynomial_decay(lr, frames(), total_frames, 0)`.
    decay, momentum, epsilon: as TensorFlow (defaults 0.9, 0.0, 1e-10; the
      reference runs .99, 0., .1 - `experiment.py:92-95`).
    group: None for one GPU, or the process group of the ranks sharing the
      batch.  Collective: every rank constructs it, with identical parameter
      shapes; parameters are made identical by a broadcast from the group's
      first rank.

  Loss convention: the reference's losses are SUMS over the batch, so with a
  sharded batch the ranks' gradients are added, not averaged.
  """

  def __init__(self, params, learning_rate, decay=0.9, momentum=0.0,
               epsilon=1e-10, use_locking=False, centered=False,
               name='RMSProp', group=None):
    del use_locking, name
    if centered:
      raise ValueError('RMSPropOptimizer: centered=True is not supported '
                       '(the reference does not use it)')
    self.params = [p for p in params]
    if not self.params:
      raise ValueError('RMSPropOptimizer: no parameters')
    dev = self.params[0].device
    for p in self.params:
      if p.device != dev or not p.is_cuda or p.dtype != torch.float32:
        raise ValueError('RMSPropOptimizer needs float32 CUDA parameters on one device')
    self.device = dev
    self.learning_rate = learning_rate
    self.decay, self.momentum, self.epsilon = float(decay), float(momentum), float(epsilon)
    self.lib = _lib.load()
    self.group = group
    self._offsets, self.n = _flatten(self.params)
    self.steps = 0

    if group is None:
      self.world, self.rank = 1, 0
      self.flat = torch.zeros(self.n, dtype=torch.float32, device=dev)
      self.flat_grad = torch.zeros(self.n, dtype=torch.float32, device=dev)
      self.lo, self.hi = 0, self.n
    else:
      import torch.distributed as dist
      import torch.distributed._symmetric_memory as symm_mem
      self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
      self.flat = symm_mem.empty(self.n, dtype=torch.float32, device=dev)
      self.flat_grad = symm_mem.empty(self.n, dtype=torch.float32, device=dev)
      self.sync = symm_mem.empty(self.lib.vtr_rmsprop_sync_bytes() // 4,
                                 dtype=torch.int32, device=dev)
      self.flat.zero_(), self.flat_grad.zero_(), self.sync.zero_()
      # a word in pinned host memory the kernel writes when it gives a step up:
      # step() looks at it without synchronising the device
      self._host_status = torch.zeros(1, dtype=torch.int32, pin_memory=True)
      self._peers = []
      for buf in (self.flat, self.flat_grad, self.sync):
        handle = symm_mem.rendezvous(buf, group)
        self._peers.append((ctypes.c_void_p * self.world)(*list(handle.buffer_ptrs)))
      lo, hi = ctypes.c_int64(0), ctypes.c_int64(0)
      _lib.check(self.lib.vtr_rmsprop_slice(self.n, self.world, self.rank,
                                            ctypes.byref(lo), ctypes.byref(hi)),
                 'vtr_rmsprop_slice')
      self.lo, self.hi = lo.value, hi.value

    with torch.no_grad():
      for p, off in zip(self.params, self._offsets):
        view = self.flat[off:off + p.numel()].view_as(p)
        view.copy_(p)
        p.data = view
        p.grad = self.flat_grad[off:off + p.numel()].view_as(p)
    if group is not None:
      import torch.distributed as dist
      dist.broadcast(self.flat, src=dist.get_global_rank(group, 0), group=group)
      torch.cuda.synchronize(dev)
      dist.barrier(group)   # every buffer is initialised before anyone steps

    # slots as tf.train.RMSPropOptimizer creates them; sharded with the slice
    span = max(self.hi - self.lo, 0)
    self.ms = torch.ones(span, dtype=torch.float32, device=dev)
    self.mom = (torch.zeros(span, dtype=torch.float32, device=dev)
                if self.momentum != 0.0 else None)

  def zero_grad(self):
    """Zeroes the flat gradient buffer (the `.grad` views stay attached)."""
    self.flat_grad.zero_()

  def current_learning_rate(self):
    lr = self.learning_rate
    return float(lr() if callable(lr) else lr)

  def step(self):
    """One `train_op` (`experiment.py:415`) on the gradients now in `.grad`."""
    for p, off in zip(self.params, self._offsets):
      if p.grad is None or p.grad.data_ptr() != self.flat_grad.data_ptr() + 4 * off:
        raise RuntimeError(
            'RMSPropOptimizer: a parameter\'s .grad no longer points into the '
            'flat gradient buffer (use optimizer.zero_grad(), not set_to_none)')
    if self.group is not None and int(self._host_status[0]) != 0:
      raise RuntimeError(
          'RMSPropOptimizer: step %d was given up because a peer did not reach '
          'the optimiser step within %d ms; parameters are no longer in sync'
          % (int(self._host_status[0]), self.lib.vtr_get_peer_timeout_ms()))
    lr = self.current_learning_rate()
    self.steps += 1
    stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
    mom = self.mom.data_ptr() if self.mom is not None else None
    with torch.cuda.device(self.device):
      if self.group is None:
        rc = self.lib.vtr_rmsprop_f32(
            self.flat.data_ptr(), self.ms.data_ptr(), mom,
            self.flat_grad.data_ptr(), self.n, lr, self.decay, self.momentum,
            self.epsilon, stream)
        _lib.check(rc, 'vtr_rmsprop_f32')
      else:
        var_peers, grad_peers, sync_peers = self._peers
        rc = self.lib.vtr_rmsprop_allreduce_f32(
            var_peers, grad_peers, sync_peers, self.ms.data_ptr(), mom, self.n,
            self.rank, self.world, self.steps, lr, self.decay, self.momentum,
            self.epsilon, self._host_status.data_ptr(), stream)
        _lib.check(rc, 'vtr_rmsprop_allreduce_f32')

  def check_peers(self):
    """Raises if a step was ever given up because a peer did not show up within
    the time-out (synchronises the device; `step()` itself checks the previous
    steps through a pinned host word, without synchronising)."""
    if self.group is not None and int(self.sync[33].item()) != 0:
      raise RuntimeError('RMSPropOptimizer: a peer did not reach the optimiser '
                         'step within %d ms; parameters are no longer in sync'
                         % self.lib.vtr_get_peer_timeout_ms())
