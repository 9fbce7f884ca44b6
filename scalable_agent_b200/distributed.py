"""Batch-sharded use of the path across the GPUs of one box.

V-trace columns are independent (`/root/reference/vtrace.py:250-262` works
elementwise over `[B]`) and the losses are plain sums over (t, b)
(`experiment.py:328,335,343`), so a data-parallel learner gives each rank its
own contiguous `[T, B/N, ...]` tensors and runs the same kernels on them.  The
only exchange is the loss scalars: they are all-gathered and summed in rank
order, so the global loss is bitwise reproducible whatever algorithm NCCL
picks.  Gradients w.r.t. logits and values are per-sample and stay local.
"""
import ctypes

import torch
import torch.distributed as dist


def shard_bounds(B, world_size, rank):
  """[lo, hi) of the batch columns rank `rank` owns (contiguous, balanced)."""
  if not 0 <= rank < world_size:
    raise ValueError('rank %d outside world of %d' % (rank, world_size))
  base, extra = divmod(B, world_size)
  lo = rank * base + min(rank, extra)
  return lo, lo + base + (1 if rank < extra else 0)


def shard_time_major(x, world_size, rank, batch_dim=1):
  """The rank's contiguous copy of its columns of a time-major tensor."""
  lo, hi = shard_bounds(x.shape[batch_dim], world_size, rank)
  return x.narrow(batch_dim, lo, hi - lo).contiguous()


def all_reduce_losses(local_losses, group=None):
  """Sum a small loss vector over ranks, deterministically.

  Args:
    local_losses: tensor [k] (e.g. the 4 loss scalars of the fused pass) on
      this rank's device.
  Returns:
    tensor [k]: sum over ranks, accumulated in float64 in rank order,
    returned in `local_losses.dtype`.  Every rank gets identical bits.
  """
  if not dist.is_available() or not dist.is_initialized():
    return local_losses.clone()
  world = dist.get_world_size(group)
  if world == 1:
    return local_losses.clone()
  flat = local_losses.detach().reshape(-1).contiguous()
  gathered = torch.empty(world * flat.numel(), dtype=flat.dtype,
                         device=flat.device)
  dist.all_gather_into_tensor(gathered, flat, group=group)
  total = torch.zeros(flat.numel(), dtype=torch.float64, device=flat.device)
  for r in range(world):            # fixed order: rank 0, 1, ...
    total += gathered[r * flat.numel():(r + 1) * flat.numel()].double()
  return total.to(local_losses.dtype).reshape(local_losses.shape)


def compute_vtrace_loss_sharded(*args, group=None, **kwargs):
  """`experiment.compute_vtrace_loss` on this rank's shard + global losses.

  Returns (local VTraceLossReturns, global_losses[4]) where global_losses =
  [total, policy_gradient, baseline, entropy] summed over ranks.  Call
  `.backward()` on the local `total_loss`: d(global total)/d(local logits) is
  exactly the local gradient, because shards share no terms.
  """
  from scalable_agent_b200 import experiment
  out = experiment.compute_vtrace_loss(*args, **kwargs)
  local = torch.stack([out.total_loss.detach(), out.policy_gradient_loss,
                       out.baseline_loss, out.entropy_loss])
  return out, all_reduce_losses(local, group=group)


class LossExchange(object):
  """Sum of the loss scalars over the ranks of one box through NVLink peer memory.

  Every rank stores its values straight into every peer's buffer (torch
  symmetric memory, P2P stores over NVLink / NVSwitch) and sums what has
  arrived in rank order: two tiny kernels of this library per step
  (`vtr_exchange_publish_f32`, `vtr_exchange_collect_f32`) instead of a
  collective library call, and identical bits on every rank.  Construct once
  (collective: every rank must), then call `all_sum` once per step on every
  rank, in the same order.
  """

  def __init__(self, device, group=None, max_values=4):
    from scalable_agent_b200 import _lib
    import torch.distributed._symmetric_memory as symm_mem
    self._lib = _lib
    self.lib = _lib.load()
    self.group = group if group is not None else dist.group.WORLD
    self.world = dist.get_world_size(self.group)
    self.rank = dist.get_rank(self.group)
    self.device = torch.device(device)
    if not 1 <= max_values <= 6:
      raise ValueError('max_values must be in 1..6')
    if self.world > 16:
      raise ValueError('LossExchange supports up to 16 ranks')
    nbytes = self.lib.vtr_exchange_buffer_bytes(self.world)
    self.buffer = symm_mem.empty(nbytes // 4, dtype=torch.float32,
                                 device=self.device)
    self.buffer.zero_()
    self.handle = symm_mem.rendezvous(self.buffer, self.group)
    ptrs = list(self.handle.buffer_ptrs)
    self._peer_ptrs = (ctypes.c_void_p * self.world)(*ptrs)
    self.seq = 0
    torch.cuda.synchronize(self.device)
    dist.barrier(self.group)   # every buffer is zeroed before anyone publishes

  def attach(self, params, sums=None, lag=0):
    """Makes the next fused call do the exchange itself: fills the exchange_*
    fields of a `_lib.FusedParams` with this exchange's buffers and the next
    sequence number.  The kernel then stores the four scalars into the peers
    from the CTA that finishes the loss reduction.  With `sums` (a float32
    CUDA tensor of 4) the same CTA also writes a sum over ranks there:

      lag=0  this step's sums: the launch ends only when the slowest rank has
             published, so every step pays the rendezvous;
      lag=1  the previous step's sums, which are normally in this rank's buffer
             already and are fetched at the start of the launch, next to the
             streaming: no rank waits for another (nothing is written on the
             first step; fetch the last step's sums with `collect()`).

    Without `sums`, follow the fused call with `collect()` on the same stream.
    Every rank must make the same choice, once per step."""
    if lag not in (0, 1):
      raise ValueError('lag must be 0 or 1')
    self.seq += 1
    params.exchange_peer_buffers = ctypes.cast(self._peer_ptrs,
                                               ctypes.POINTER(ctypes.c_void_p))
    params.exchange_rank, params.exchange_world = self.rank, self.world
    params.exchange_seq = self.seq
    params.exchange_lag = lag
    if sums is not None:
      if (sums.dtype != torch.float32 or sums.numel() < 4 or
          not sums.is_contiguous() or sums.device != self.device):
        raise ValueError('sums must be a contiguous float32[4] tensor on %s' % self.device)
      params.exchange_sums = sums.data_ptr()
    else:
      params.exchange_sums = None
    return self.seq

  def collect(self, n=4):
    """Sum over ranks of what the ranks published for the current step (waits
    on the device until every rank has)."""
    out = torch.empty(n, dtype=torch.float32, device=self.device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
    with torch.cuda.device(self.device):
      rc = self.lib.vtr_exchange_collect_f32(
          self.buffer.data_ptr(), n, self.world, self.seq, out.data_ptr(),
          stream)
      self._lib.check(rc, 'vtr_exchange_collect_f32')
    return out

  def all_sum(self, values):
    """values: float32 tensor [n], n <= 6, on this rank's device -> sum over ranks."""
    flat = values.detach().reshape(-1)
    if flat.dtype != torch.float32 or not flat.is_contiguous():
      flat = flat.to(torch.float32).contiguous()
    n = flat.numel()
    out = torch.empty(n, dtype=torch.float32, device=self.device)
    self.seq += 1
    stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
    with torch.cuda.device(self.device):
      rc = self.lib.vtr_exchange_publish_f32(
          flat.data_ptr(), n, self._peer_ptrs, self.rank, self.world, self.seq,
          stream)
      self._lib.check(rc, 'vtr_exchange_publish_f32')
      rc = self.lib.vtr_exchange_collect_f32(
          self.buffer.data_ptr(), n, self.world, self.seq, out.data_ptr(),
          stream)
      self._lib.check(rc, 'vtr_exchange_collect_f32')
    return out.reshape(values.shape)
