"""The NVLink peer-memory loss exchange (vtr_exchange_*).

The two-rank tests need two CUDA devices and are skipped on a one-GPU box (the
driver's `-m gpu` run); run them with
  gpurun --gpus 2 -- python -m pytest tests/test_exchange_gpu.py -m gpu
The world-of-one tests run anywhere: the "peer" is the rank's own buffer.
"""
import ctypes
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

TWO_GPU_REASON = (
    'needs 2 GPUs: two ranks on ONE device would be two processes whose kernels wait for each '
    "other's stores, which a single GPU may time-slice instead of running side by side (a wait "
    'could then starve the rank it waits for); bench.py exchange_check / optim_check carry the '
    'proof on every multi-GPU run')



def _free_port():
  with socket.socket() as s:
    s.bind(('127.0.0.1', 0))
    return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  torch.cuda.set_device(rank)
  device = torch.device('cuda', rank)
  dist.init_process_group('nccl', rank=rank, world_size=world, device_id=device)
  try:
    from scalable_agent_b200.distributed import LossExchange
    ex = LossExchange(device)
    results = []
    for step in range(1, 6):
      mine = torch.tensor([rank + 0.25 * step, -3.0 * rank, 1e-3 * step, 7.0],
                          device=device)
      results.append(ex.all_sum(mine).cpu())
    torch.save(torch.stack(results), os.path.join(out_dir, 'r%d.pt' % rank))
  finally:
    dist.destroy_process_group()


def test_loss_exchange_two_ranks(tmp_path):
  if torch.cuda.device_count() < 2:
    pytest.skip(TWO_GPU_REASON)
  world = 2
  mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world,
           join=True)
  r0 = torch.load(tmp_path / 'r0.pt')
  r1 = torch.load(tmp_path / 'r1.pt')
  assert torch.equal(r0, r1), 'every rank must hold the same bits'
  for i, step in enumerate(range(1, 6)):
    want = torch.tensor([(0 + 0.25 * step) + (1 + 0.25 * step), -3.0,
                         2e-3 * step, 14.0])
    torch.testing.assert_close(r0[i], want, rtol=1e-6, atol=1e-6)


# ---------------------------------------------------------------------------
# The fused pass publishing its own losses (exchange_* fields of vtr_fused_params)
# ---------------------------------------------------------------------------
KEYS = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
        'discounts', 'rewards', 'values', 'bootstrap_value')
# (T, B, A): the per-SM kernel, the cluster kernel, the item kernel and the
# 32-column-tile kernels (here: more steps than a cluster holds); each publishes
# from the CTA that finishes the loss reduction
FUSED_SHAPES = [(32, 64 * 148, 6), (7, 50, 5), (9, 8, 4), (300, 40, 5)]


def _fused_call(_lib, lib, T, B, A, seed, device):
  """Inputs on `device`, params with every output, and the workspace."""
  from oracle import vtrace_oracle as oracle
  inp = oracle.synthetic_inputs(T, B, A, seed=seed)
  keep = {k: torch.from_numpy(inp[k]).to(device) for k in KEYS}
  f32 = dict(dtype=torch.float32, device=device)
  keep.update(vs=torch.empty(T, B, **f32), pg_advantages=torch.empty(T, B, **f32),
              d_target_logits=torch.empty(T, B, A, **f32),
              d_values=torch.empty(T, B, **f32), losses=torch.zeros(4, **f32))
  p = _lib.FusedParams()
  p.T, p.B, p.A = T, B, A
  p.clip_rho_threshold = p.clip_pg_rho_threshold = 1.0
  p.baseline_cost, p.entropy_cost = 0.5, 0.00025
  for k, v in keep.items():
    setattr(p, k, v.data_ptr())
  nbytes = lib.vtr_fused_workspace_bytes(T, B, A)
  keep['ws'] = torch.empty(nbytes // 8, dtype=torch.float64, device=device)
  return p, keep, nbytes


@pytest.mark.parametrize('two_launches', [False, True])
def test_fused_pass_publishes_its_losses_world_of_one(monkeypatch, two_launches):
  """two_launches: the final reduction left to K3 (as under stream capture),
  so the library appends the publish / collect launches itself."""
  assert torch.cuda.is_available()
  if two_launches:
    monkeypatch.setenv('VTR_NO_INKERNEL_FINISH', '1')
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  device = torch.device('cuda', 0)
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  buf = torch.zeros(lib.vtr_exchange_buffer_bytes(1) // 4, device=device)
  peers = (ctypes.c_void_p * 1)(buf.data_ptr())
  seq = 0
  for T, B, A in FUSED_SHAPES * 2:        # both parities of the slot for each shape
    p, keep, nbytes = _fused_call(_lib, lib, T, B, A, seed=B + seq, device=device)
    plain = torch.zeros(4, device=device)
    p.losses = plain.data_ptr()
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
    assert rc == 0, _lib.last_error()
    seq += 1
    p.losses = keep['losses'].data_ptr()
    p.exchange_peer_buffers = ctypes.cast(peers, ctypes.POINTER(ctypes.c_void_p))
    p.exchange_rank, p.exchange_world, p.exchange_seq = 0, 1, seq
    count = ctypes.c_int(0)
    text = ctypes.create_string_buffer(512)
    assert lib.vtr_fused_describe(ctypes.byref(p), 1, ctypes.byref(count), text, 512) == 0
    inside = b'peer publish' in text.value
    assert inside == (not two_launches or B >= 64 * 148), text.value
    assert inside or b'exchange_publish_kernel' in text.value
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
    assert rc == 0, _lib.last_error()
    got = torch.empty(4, device=device)
    rc = lib.vtr_exchange_collect_f32(buf.data_ptr(), 4, 1, seq, got.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(got, plain), (got, plain)
    assert torch.equal(keep['losses'], plain)
    # and with the collect done by the pass as well
    seq += 1
    sums = torch.zeros(4, device=device)
    p.exchange_seq, p.exchange_sums = seq, sums.data_ptr()
    assert lib.vtr_fused_describe(ctypes.byref(p), 1, ctypes.byref(count), text, 512) == 0
    assert (b'peer publish + collect' in text.value) == inside, text.value
    assert inside or b'exchange_collect_kernel' in text.value
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(sums, plain), (sums, plain)
    # lag 1: the pass hands out the previous step's sums and publishes its own
    seq += 1
    p.exchange_seq, p.exchange_lag = seq, 1
    sums.fill_(-1.0)
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
    assert rc == 0, _lib.last_error()
    rc = lib.vtr_exchange_collect_f32(buf.data_ptr(), 4, 1, seq, got.data_ptr(), stream)
    assert rc == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(sums, plain) and torch.equal(got, plain), (sums, got, plain)
    p.exchange_lag = 0


def test_fused_pass_lagged_sums_first_step_writes_nothing():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  device = torch.device('cuda', 0)
  stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
  buf = torch.zeros(lib.vtr_exchange_buffer_bytes(1) // 4, device=device)
  peers = (ctypes.c_void_p * 1)(buf.data_ptr())
  p, keep, nbytes = _fused_call(_lib, lib, 7, 50, 5, seed=3, device=device)
  sums = torch.full((4,), -7.0, device=device)
  p.exchange_peer_buffers = ctypes.cast(peers, ctypes.POINTER(ctypes.c_void_p))
  p.exchange_rank, p.exchange_world, p.exchange_seq, p.exchange_lag = 0, 1, 1, 1
  p.exchange_sums = sums.data_ptr()
  rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
  assert rc == 0, _lib.last_error()
  torch.cuda.synchronize()
  assert torch.equal(sums, torch.full((4,), -7.0, device=device))
  p.exchange_lag = 2
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream) == -1


def test_fused_pass_rejects_bad_exchange_fields():
  assert torch.cuda.is_available()
  from scalable_agent_b200 import _lib
  lib = _lib.load()
  device = torch.device('cuda', 0)
  p, keep, nbytes = _fused_call(_lib, lib, 7, 50, 5, seed=3, device=device)
  buf = torch.zeros(16, device=device)
  peers = (ctypes.c_void_p * 1)(buf.data_ptr())
  p.exchange_peer_buffers = ctypes.cast(peers, ctypes.POINTER(ctypes.c_void_p))
  for rank, world, seq in [(0, 1, 0), (1, 1, 1), (0, 17, 1), (-1, 1, 1)]:
    p.exchange_rank, p.exchange_world, p.exchange_seq = rank, world, seq
    rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, None)
    assert rc == -1, (rank, world, seq)
  p.exchange_rank, p.exchange_world, p.exchange_seq = 0, 1, 1
  p.losses = None
  assert lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, None) == -1


def _fused_worker(rank, world, port, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  torch.cuda.set_device(rank)
  device = torch.device('cuda', rank)
  dist.init_process_group('nccl', rank=rank, world_size=world, device_id=device)
  try:
    from scalable_agent_b200 import _lib
    from scalable_agent_b200.distributed import LossExchange
    lib = _lib.load()
    ex = LossExchange(device)
    stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    sums, locals_ = [], []
    for step, (T, B, A) in enumerate(FUSED_SHAPES * 3):
      p, keep, nbytes = _fused_call(_lib, lib, T, B, A, seed=100 * rank + step, device=device)
      inside = torch.zeros(4, device=device) if step % 3 else None
      ex.attach(p, inside)          # every rank: collect inside on the same steps
      rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
      assert rc == 0, _lib.last_error()
      sums.append((ex.collect() if inside is None else inside).cpu())
      locals_.append(keep['losses'].cpu())
    # lagged: step k hands out the sums of step k-1; rank 1 is held back now
    # and then so that the ranks drift
    lagged = []
    prev = torch.zeros(4, device=device)
    for step, (T, B, A) in enumerate(FUSED_SHAPES * 3):
      p, keep, nbytes = _fused_call(_lib, lib, T, B, A, seed=7000 + 100 * rank + step, device=device)
      if rank == 1 and step % 4 == 1:
        torch.cuda._sleep(3000000)
      ex.attach(p, prev, lag=1)
      rc = lib.vtr_from_logits_fused_f32(ctypes.byref(p), keep['ws'].data_ptr(), nbytes, stream)
      assert rc == 0, _lib.last_error()
      if step > 0:
        lagged.append(prev.cpu())     # sums of step - 1
      locals_.append(keep['losses'].cpu())
    lagged.append(ex.collect().cpu())  # the last step's
    sums.extend(lagged)
    torch.save((torch.stack(sums), torch.stack(locals_)),
               os.path.join(out_dir, 'f%d.pt' % rank))
  finally:
    dist.destroy_process_group()


def test_fused_pass_publishes_to_peers_two_ranks(tmp_path):
  if torch.cuda.device_count() < 2:
    pytest.skip(TWO_GPU_REASON)
  world = 2
  mp.spawn(_fused_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world,
           join=True)
  s0, l0 = torch.load(tmp_path / 'f0.pt')
  s1, l1 = torch.load(tmp_path / 'f1.pt')
  assert torch.equal(s0, s1), 'every rank must hold the same bits'
  want = (l0.double() + l1.double()).float()      # rank order, float64, as the collect does
  assert torch.equal(s0, want), (s0, want)
  assert np.all(np.isfinite(s0.numpy())) and float(s0.abs().min()) > 0
