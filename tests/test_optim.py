"""The optimiser step after the path (experiment.py:410-415): polynomial
learning-rate decay + RMSProp, single GPU and sharded over NVLink peers.

Oracle: oracle/rmsprop_oracle.py (float64 restatement of TensorFlow 1.9.0's
published ApplyRMSProp / polynomial_decay; parity unpinned - the reference has
no test for its optimiser).  Tolerance: 1e-5 * max(1, |ref|), as everywhere.
"""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch

from oracle import rmsprop_oracle as oracle
from tests.conftest import assert_close

# experiment.py:92-95
LR, DECAY, MOMENTUM, EPSILON = 0.00048, 0.99, 0.0, 0.1

TWO_GPU_REASON = (
    'needs 2 GPUs: two ranks on ONE device would be two processes whose kernels wait for each '
    "other's stores, which a single GPU may time-slice instead of running side by side (a wait "
    'could then starve the rank it waits for); bench.py exchange_check / optim_check carry the '
    'proof on every multi-GPU run')



# ---------------------------------------------------------------------------
# CPU: the oracle itself and the host arithmetic of the library
# ---------------------------------------------------------------------------
def test_oracle_polynomial_decay_is_linear_to_zero():
  # experiment.py:411-412: lr0 -> 0 over total_environment_frames, then stays
  assert oracle.polynomial_decay(LR, 0, 1e9, 0) == LR
  assert oracle.polynomial_decay(LR, 5e8, 1e9, 0) == pytest.approx(LR / 2, rel=1e-15)
  assert oracle.polynomial_decay(LR, 1e9, 1e9, 0) == 0.0
  assert oracle.polynomial_decay(LR, 3e9, 1e9, 0) == 0.0
  assert oracle.polynomial_decay(1.0, 50, 100, 0.2, power=2.0) == pytest.approx(0.8 * 0.25 + 0.2)


def test_oracle_rmsprop_first_step_by_hand():
  # ms0 = 1: ms = .99 + .01 g^2 ; step = lr g / sqrt(ms + .1)
  var, g = np.array([1.0, -2.0]), np.array([3.0, 0.5])
  ms0, mom0 = oracle.rmsprop_slots(var)
  assert np.all(ms0 == 1.0) and np.all(mom0 == 0.0)
  v, ms, mom = oracle.rmsprop_step(var, ms0, mom0, g, LR, DECAY, MOMENTUM, EPSILON)
  want_ms = 0.99 + 0.01 * g * g
  np.testing.assert_allclose(ms, want_ms, rtol=1e-15)
  np.testing.assert_allclose(v, var - LR * g / np.sqrt(want_ms + 0.1), rtol=1e-15)
  # momentum carries the previous update
  v2, _, mom2 = oracle.rmsprop_step(v, ms, mom, g, LR, DECAY, 0.5, EPSILON)
  ms2 = 0.99 * want_ms + 0.01 * g * g
  np.testing.assert_allclose(mom2, 0.5 * mom + LR * g / np.sqrt(ms2 + 0.1), rtol=1e-15)
  np.testing.assert_allclose(v2, v - mom2, rtol=1e-15)


def test_oracle_matches_torch_float64_restatement():
  """Independent restatement with torch ops (TF's `ms += (g^2 - ms)(1 - rho)` form)."""
  rng = np.random.RandomState(5)
  var, g = rng.randn(1000), rng.randn(1000) * 3
  ms, mom = oracle.rmsprop_slots(var)
  tv, tg = torch.from_numpy(var.copy()), torch.from_numpy(g)
  tms, tmom = torch.ones_like(tv), torch.zeros_like(tv)
  for _ in range(4):
    var, ms, mom = oracle.rmsprop_step(var, ms, mom, g, 0.01, 0.9, 0.3, 1e-3)
    tms += (tg * tg - tms) * (1 - 0.9)
    tmom = tmom * 0.3 + (tg * 0.01) / (tms + 1e-3).sqrt()
    tv -= tmom
  np.testing.assert_allclose(var, tv.numpy(), rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize('n,world', [(0, 1), (1, 1), (7, 2), (8, 2), (1000, 8), (1001, 8),
                                     (3, 4), (1 << 20, 3), (12345, 16)])
def test_slices_partition_the_parameters(n, world):
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  covered = 0
  for rank in range(world):
    lo, hi = ctypes.c_int64(-1), ctypes.c_int64(-1)
    assert lib.vtr_rmsprop_slice(n, world, rank, ctypes.byref(lo), ctypes.byref(hi)) == 0
    assert (lo.value, hi.value) == oracle.slice_bounds(n, world, rank)
    assert lo.value == covered and lo.value % 4 == 0 or lo.value == n
    assert hi.value >= lo.value
    covered = hi.value
  assert covered == n
  lo, hi = ctypes.c_int64(0), ctypes.c_int64(0)
  assert lib.vtr_rmsprop_slice(10, 2, 2, ctypes.byref(lo), ctypes.byref(hi)) == -1
  assert lib.vtr_rmsprop_slice(10, 17, 0, ctypes.byref(lo), ctypes.byref(hi)) == -1


def test_library_polynomial_decay_matches_oracle():
  from scalable_agent_b200 import optim
  for step in [0, 1, 12345678, int(5e8), int(1e9), int(2e9)]:
    got = optim.polynomial_decay(LR, step, int(1e9), 0)
    assert got == pytest.approx(oracle.polynomial_decay(LR, step, 1e9, 0), rel=1e-15, abs=0)
  assert optim.polynomial_decay(1.0, 50, 100, 0.2, power=2.0) == pytest.approx(0.4)
  # tf.train.polynomial_decay's default end_learning_rate is 0.0001
  assert optim.polynomial_decay(1.0, 100, 100) == pytest.approx(0.0001)
  with pytest.raises(ValueError):
    optim.polynomial_decay(1.0, 1, 10, cycle=True)


# ---------------------------------------------------------------------------
# GPU: the kernels, through the C ABI
# ---------------------------------------------------------------------------
def _run_steps(lib, n, steps, lr, decay, momentum, eps, seed, with_mom_slot):
  rng = np.random.RandomState(seed)
  var = rng.randn(n).astype(np.float32)
  grads = [(rng.randn(n) * 10 ** rng.uniform(-3, 1)).astype(np.float32) for _ in range(steps)]
  d_var = torch.from_numpy(var.copy()).cuda()
  d_ms = torch.ones(n, device='cuda')
  d_mom = torch.zeros(n, device='cuda') if with_mom_slot else None
  o_var = var.astype(np.float64)
  o_ms, o_mom = oracle.rmsprop_slots(o_var)
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  for g in grads:
    d_g = torch.from_numpy(g).cuda()
    rc = lib.vtr_rmsprop_f32(d_var.data_ptr(), d_ms.data_ptr(),
                             d_mom.data_ptr() if d_mom is not None else None,
                             d_g.data_ptr(), n, lr, decay, momentum, eps, stream)
    assert rc == 0
    o_var, o_ms, o_mom = oracle.rmsprop_step(o_var, o_ms, o_mom, g, lr, decay, momentum, eps)
  torch.cuda.synchronize()
  return d_var, d_ms, d_mom, o_var, o_ms, o_mom


@pytest.mark.gpu
@pytest.mark.parametrize('n', [1, 3, 4, 5, 1023, 4096, 100003, 1 << 21])
@pytest.mark.parametrize('momentum', [0.0, 0.9])
def test_rmsprop_matches_oracle(n, momentum):
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  d_var, d_ms, d_mom, o_var, o_ms, o_mom = _run_steps(
      lib, n, 5, 0.01, DECAY, momentum, EPSILON, seed=n, with_mom_slot=True)
  assert_close(d_var.cpu().numpy(), o_var, 1e-5, 'var')
  assert_close(d_ms.cpu().numpy(), o_ms, 1e-5, 'ms')
  assert_close(d_mom.cpu().numpy(), o_mom, 1e-5, 'mom')


@pytest.mark.gpu
def test_rmsprop_without_momentum_slot_and_argument_errors():
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  n = 5001
  with_slot = _run_steps(lib, n, 3, LR, DECAY, 0.0, EPSILON, seed=1, with_mom_slot=True)
  no_slot = _run_steps(lib, n, 3, LR, DECAY, 0.0, EPSILON, seed=1, with_mom_slot=False)
  assert torch.equal(with_slot[0], no_slot[0]) and torch.equal(with_slot[1], no_slot[1])
  assert_close(no_slot[0].cpu().numpy(), no_slot[3], 1e-5, 'var')
  x = torch.ones(8, device='cuda')
  p = x.data_ptr()
  assert lib.vtr_rmsprop_f32(p, p, None, p, 8, LR, DECAY, 0.5, EPSILON, None) == -1   # slot needed
  assert lib.vtr_rmsprop_f32(p, p, p, p, 8, LR, 1.5, 0.0, EPSILON, None) == -1        # decay
  assert lib.vtr_rmsprop_f32(p + 4, p, p, p, 4, LR, DECAY, 0.0, EPSILON, None) == -4  # alignment
  assert lib.vtr_rmsprop_f32(None, p, p, p, 8, LR, DECAY, 0.0, EPSILON, None) == -1
  assert lib.vtr_rmsprop_f32(p, p, p, p, 0, LR, DECAY, 0.0, EPSILON, None) == 0


@pytest.mark.gpu
def test_optimizer_trains_a_torch_module_like_the_oracle():
  """RMSPropOptimizer over nn.Module parameters with autograd gradients and the
  reference's schedule (experiment.py:410-415)."""
  from scalable_agent_b200 import optim
  torch.manual_seed(0)
  net = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3)).cuda()
  shadow = [p.detach().cpu().double().numpy().copy() for p in net.parameters()]
  slots = [oracle.rmsprop_slots(s) for s in shadow]
  frames = [0]
  opt = optim.RMSPropOptimizer(
      net.parameters(), lambda: optim.polynomial_decay(0.05, frames[0], 1000, 0),
      DECAY, 0.3, EPSILON)
  x = torch.randn(11, 7, device='cuda')
  for step in range(6):
    opt.zero_grad()
    loss = net(x).square().sum()          # a sum, like the reference's losses
    loss.backward()
    grads = [p.grad.detach().cpu().double().numpy().copy() for p in net.parameters()]
    lr = oracle.polynomial_decay(0.05, frames[0], 1000, 0)
    opt.step()
    for i, g in enumerate(grads):
      v, ms, mom = oracle.rmsprop_step(shadow[i], slots[i][0], slots[i][1], g, lr, DECAY, 0.3,
                                       EPSILON)
      shadow[i], slots[i] = v, (ms, mom)
    frames[0] += 150
    # the oracle follows the kernel's trajectory: compare, then re-base it on the
    # device values so that float32 drift of the network does not accumulate
    for i, p in enumerate(net.parameters()):
      assert_close(p.detach().cpu().numpy(), shadow[i], 1e-5, 'step %d param %d' % (step, i))
      shadow[i] = p.detach().cpu().double().numpy().copy()
  assert opt.current_learning_rate() == pytest.approx(0.05 * (1 - 900 / 1000))
  # parameters are views of one flat buffer and stay usable by the module
  assert all(p.data_ptr() >= opt.flat.data_ptr() for p in net.parameters())
  with pytest.raises(RuntimeError):
    next(net.parameters()).grad = None
    opt.step()


def _sharded_world_of_one(lib, n, steps, momentum):
  """vtr_rmsprop_allreduce_f32 with this GPU as its own only peer."""
  rng = np.random.RandomState(n)
  var = rng.randn(n).astype(np.float32)
  d_var = torch.from_numpy(var.copy()).cuda()
  d_grad = torch.zeros(n, device='cuda')
  sync = torch.zeros(lib.vtr_rmsprop_sync_bytes() // 4, dtype=torch.int32, device='cuda')
  d_ms = torch.ones(n, device='cuda')
  d_mom = torch.zeros(n, device='cuda') if momentum else None
  peers = [(ctypes.c_void_p * 1)(t.data_ptr()) for t in (d_var, d_grad, sync)]
  o_var = var.astype(np.float64)
  o_ms, o_mom = oracle.rmsprop_slots(o_var)
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  for seq in range(1, steps + 1):
    g = rng.randn(n).astype(np.float32)
    d_grad.copy_(torch.from_numpy(g))
    rc = lib.vtr_rmsprop_allreduce_f32(
        peers[0], peers[1], peers[2], d_ms.data_ptr(),
        d_mom.data_ptr() if d_mom is not None else None, n, 0, 1, seq,
        0.01, DECAY, momentum, EPSILON, None, stream)
    assert rc == 0
    o_var, o_ms, o_mom = oracle.rmsprop_step(o_var, o_ms, o_mom, g, 0.01, DECAY, momentum, EPSILON)
  torch.cuda.synchronize()
  assert int(sync[33].item()) == 0 and int(sync[32].item()) == 0
  return d_var, d_ms, o_var, o_ms


@pytest.mark.gpu
@pytest.mark.parametrize('n', [1, 6, 4096, 99999, 1 << 20])
@pytest.mark.parametrize('momentum', [0.0, 0.5])
def test_sharded_step_world_of_one_matches_oracle(n, momentum):
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  d_var, d_ms, o_var, o_ms = _sharded_world_of_one(lib, n, 4, momentum)
  assert_close(d_var.cpu().numpy(), o_var, 1e-5, 'var')
  assert_close(d_ms.cpu().numpy(), o_ms, 1e-5, 'ms')


# ---------------------------------------------------------------------------
# GPU x2: the sharded step between real peers
# ---------------------------------------------------------------------------
def _free_port():
  with socket.socket() as s:
    s.bind(('127.0.0.1', 0))
    return s.getsockname()[1]


SHAPES = [(33, 17), (17,), (5, 4, 3), (1,)]      # 561 + 17 + 60 + 1 parameters, odd sizes


def _worker(rank, world, port, out_dir):
  import torch.distributed as dist
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  torch.cuda.set_device(rank)
  device = torch.device('cuda', rank)
  dist.init_process_group('nccl', rank=rank, world_size=world, device_id=device)
  try:
    from scalable_agent_b200 import optim
    gen = torch.Generator().manual_seed(100 + rank)      # ranks start apart on purpose
    params = [torch.nn.Parameter(torch.randn(*s, generator=gen).to(device)) for s in SHAPES]
    opt = optim.RMSPropOptimizer(params, 0.02, DECAY, 0.4, EPSILON, group=dist.group.WORLD)
    start = [p.detach().cpu().clone() for p in params]
    grads = []
    for step in range(5):
      opt.zero_grad()
      mine = []
      for p in params:
        g = torch.randn(*p.shape, generator=gen) * (1 + rank)
        p.grad.copy_(g.to(device))
        mine.append(g)
      grads.append(mine)
      opt.step()
    torch.cuda.synchronize()
    opt.check_peers()
    torch.save({'start': start, 'grads': grads,
                'final': [p.detach().cpu() for p in params],
                'slice': (opt.lo, opt.hi), 'ms': opt.ms.cpu()},
               os.path.join(out_dir, 'o%d.pt' % rank))
  finally:
    dist.destroy_process_group()


def _late_worker(rank, world, port, out_dir):
  """Rank 1 reaches its second optimiser step long after the peer time-out."""
  import torch.distributed as dist
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  torch.cuda.set_device(rank)
  device = torch.device('cuda', rank)
  dist.init_process_group('nccl', rank=rank, world_size=world, device_id=device)
  try:
    from scalable_agent_b200 import _lib, optim
    lib = _lib.load()
    assert lib.vtr_get_peer_timeout_ms() == 20000          # the default
    lib.vtr_set_peer_timeout_ms(50)
    params = [torch.nn.Parameter(torch.zeros(4096, device=device))]
    opt = optim.RMSPropOptimizer(params, 0.02, DECAY, 0.0, EPSILON, group=dist.group.WORLD)
    params[0].grad.fill_(1.0 + rank)
    opt.step()                                             # step 1: everybody on time
    torch.cuda.synchronize()
    opt.check_peers()
    after_one = params[0].detach().cpu().clone()
    dist.barrier()
    if rank == 1:
      torch.cuda._sleep(int(0.4 * 1.9e9))                  # ~0.4 s on the GPU, 8x the time-out
    opt.step()                                             # step 2: rank 0 gives up waiting
    torch.cuda.synchronize()
    flagged = int(opt.sync[33].item())
    host_word = int(opt._host_status[0])
    raised = False
    try:
      opt.step()                                           # refuses: the host word is set
    except RuntimeError as e:
      raised = 'given up' in str(e)
    torch.save({'after_one': after_one, 'final': params[0].detach().cpu(),
                'flagged': flagged, 'host_word': host_word, 'raised': raised},
               os.path.join(out_dir, 'late%d.pt' % rank))
  finally:
    dist.destroy_process_group()


@pytest.mark.gpu
def test_sharded_optimizer_flags_a_late_peer_on_every_rank(tmp_path):
  """A rank that arrives after the time-out must not leave a silently
  half-updated model: both ranks flag the step (device word, pinned host
  word, exception on the next step), and neither slice is half-written."""
  if torch.cuda.device_count() < 2:
    pytest.skip(TWO_GPU_REASON)
  import torch.multiprocessing as mp
  world = 2
  mp.spawn(_late_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
  r = [torch.load(tmp_path / ('late%d.pt' % k)) for k in range(world)]
  assert torch.equal(r[0]['after_one'], r[1]['after_one'])
  for k in range(world):
    assert r[k]['flagged'] == 1 and r[k]['host_word'] == 2 and r[k]['raised'], (k, r[k])
    # each rank's slice is either untouched or fully updated
    for lo, hi in ((0, 2048), (2048, 4096)):
      moved = (r[k]['final'][lo:hi] != r[k]['after_one'][lo:hi])
      assert bool(moved.all()) or not bool(moved.any()), (k, lo)


@pytest.mark.gpu
def test_sharded_optimizer_two_ranks(tmp_path):
  if torch.cuda.device_count() < 2:
    pytest.skip(TWO_GPU_REASON)
  import torch.multiprocessing as mp
  world = 2
  mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
  r = [torch.load(tmp_path / ('o%d.pt' % k)) for k in range(world)]
  for a, b in zip(r[0]['start'], r[1]['start']):
    assert torch.equal(a, b), 'parameters are broadcast from rank 0 at construction'
  for a, b in zip(r[0]['final'], r[1]['final']):
    assert torch.equal(a, b), 'every rank must hold the same bits'
  assert r[0]['slice'][0] == 0 and r[0]['slice'][1] == r[1]['slice'][0]
  for i, start in enumerate(r[0]['start']):
    var = start.double().numpy()
    ms, mom = oracle.rmsprop_slots(var)
    for step in range(5):
      var, ms, mom = oracle.data_parallel_step(
          var, ms, mom, [r[k]['grads'][step][i].numpy() for k in range(world)],
          0.02, DECAY, 0.4, EPSILON)
    assert_close(r[0]['final'][i].numpy(), var, 1e-5, 'param %d' % i)
