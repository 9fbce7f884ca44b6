#!/usr/bin/env python
"""Benchmark of the hot path: fused V-trace + IMPALA loss, forward+backward.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

Metric (BASELINE.json): timesteps/s (T x B per call) of the fused pass and the
HBM bandwidth it achieves.  Workload: the configuration the metric's target is
quoted on, T=100, B=32768, A=18 per GPU (weak scaling: every rank holds its
own B columns, only the loss scalars cross NVLink).  A "step" is one call of
`vtr_from_logits_fused_f32` (K1 column kernel + K3 loss reduction) with every
output materialised, plus, for N>1, the all-gather of the loss scalars.

Prints ONE JSON line on rank 0.  See DESIGN.md section "Measurement".
"""
import argparse
import ctypes
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

T_DEFAULT, B_DEFAULT, A_DEFAULT = 100, 32768, 18
BASELINE_COST, ENTROPY_COST = 0.5, 0.00025   # experiment.py:70-71
METRIC = 'vtrace_impala_loss_fwd_bwd_timesteps_per_sec'
UNIT = 'timesteps/s'
FUSED_KEYS = ('behaviour_policy_logits', 'target_policy_logits', 'actions',
              'discounts', 'rewards', 'values', 'bootstrap_value')


def algorithmic_bytes(T, B, A):
  """(12A + 40) bytes per timestep + 4B for the bootstrap."""
  return T * B * (12 * A + 40) + 4 * B


def workload_config(T, B, A, world, scaling):
  """The `config` object; the same keys and values from both arms."""
  return {
      'workload': 'fused V-trace + IMPALA loss fwd+bwd, T=%d B=%d A=%d per GPU, '
                  'all 5 V-trace fields + d_target_logits + d_values + 4 loss '
                  'scalars' % (T, B, A),
      'T': T, 'B_per_gpu': B, 'A': A, 'global_B': B * world,
      'scaling': scaling, 'clip_rho': 1.0, 'clip_pg_rho': 1.0,
      'baseline_cost': BASELINE_COST, 'entropy_cost': ENTROPY_COST}


def measured_peak_gbs():
  path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
  try:
    with open(path) as f:
      return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
  except Exception:  # pylint: disable=broad-except
    return 6650.0, 'fallback (B200_PROFILING.md)'


def synthetic_on_device(T, B, A, seed, device):
  """SURVEY.md 8(d) recipe, generated on the device."""
  import torch
  g = torch.Generator(device=device).manual_seed(seed)
  rn = lambda *s: torch.randn(*s, generator=g, device=device)
  behaviour = rn(T, B, A)
  target = behaviour + 0.5 * rn(T, B, A)
  actions = torch.randint(0, A, (T, B), generator=g, device=device,
                          dtype=torch.int32)
  done = torch.rand(T, B, generator=g, device=device) < 0.01
  rewards = (2.0 * rn(T, B)).clamp_(-1.0, 1.0)      # abs_one, experiment.py:378
  discounts = (~done).float() * 0.99                # experiment.py:384
  values, bootstrap = rn(T, B), rn(B)
  return dict(behaviour_policy_logits=behaviour, target_policy_logits=target,
              actions=actions, discounts=discounts, rewards=rewards,
              values=values, bootstrap_value=bootstrap)


class ClockSampler(threading.Thread):
  """Samples SM clock and throttle reasons of one GPU through NVML."""

  REASONS = {
      0x8: 'hw_slowdown', 0x40: 'hw_thermal_slowdown',
      0x20: 'sw_thermal_slowdown', 0x4: 'sw_power_cap',
      0x80: 'hw_power_brake_slowdown', 0x2: 'applications_clocks_setting',
      0x100: 'display_clock_setting', 0x10: 'sync_boost',
  }

  def __init__(self, index, period=0.002):
    super().__init__(daemon=True)
    self.period, self.samples, self.reason_bits = period, [], 0
    self._stop_event = threading.Event()
    self.sm_max = None
    self.ok = False
    try:
      import pynvml
      pynvml.nvmlInit()
      self.nv = pynvml
      self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
      self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
      self.ok = True
    except Exception:  # pylint: disable=broad-except
      self.ok = False

  def run(self):
    if not self.ok:
      return
    while not self._stop_event.is_set():
      try:
        self.samples.append(
            self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
        self.reason_bits |= int(
            self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
      except Exception:  # pylint: disable=broad-except
        pass
      time.sleep(self.period)

  def stop(self):
    self._stop_event.set()
    if self.is_alive():
      self.join()

  def summary(self):
    reasons = [name for bit, name in self.REASONS.items()
               if self.reason_bits & bit]
    return {
        'sm_mhz': statistics.median(self.samples) if self.samples else None,
        'sm_max_mhz': self.sm_max, 'reasons': reasons,
        'samples': len(self.samples)}


def physical_gpu_index(local_index):
  vis = os.environ.get('CUDA_VISIBLE_DEVICES')
  if vis:
    try:
      return int(vis.split(',')[local_index])
    except (ValueError, IndexError):
      pass
  return local_index


# ---------------------------------------------------------------------------
# CPU legs (the oracle's float32 op-for-op port; see oracle/cpu_port.py)
# ---------------------------------------------------------------------------
def time_cpu_port(T, B, A, steps, warmup, seed=1):
  import torch
  from oracle import cpu_port
  from oracle import vtrace_oracle as oracle
  cores = len(os.sched_getaffinity(0))
  torch.set_num_threads(cores)
  inp = oracle.synthetic_inputs(T, B, A, seed=seed)
  args = [torch.from_numpy(inp[k]) for k in FUSED_KEYS]
  for _ in range(warmup):
    cpu_port.learner_loss(*args, baseline_cost=BASELINE_COST,
                          entropy_cost=ENTROPY_COST)
  t0 = time.perf_counter()
  for _ in range(steps):
    cpu_port.learner_loss(*args, baseline_cost=BASELINE_COST,
                          entropy_cost=ENTROPY_COST)
  dt = time.perf_counter() - t0
  return T * B * steps / dt, dt / steps, cores


def cpu_model():
  try:
    with open('/proc/cpuinfo') as f:
      for line in f:
        if line.startswith('model name'):
          return line.split(':', 1)[1].strip()
  except OSError:
    pass
  return 'unknown'


def pick_cpu_sample(T, A, steps, warmup, budget_s):
  """Largest B (<= B_DEFAULT) whose steps+warmup fit in budget_s."""
  rate, _, _ = time_cpu_port(T, 2048, A, 3, 2)
  for B in (32768, 16384, 8192, 4096, 2048, 1024, 512, 256):
    if T * B * (steps + warmup) / rate <= budget_s:
      return B
  return 128


def run_reference(args):
  """--impl reference: the reference's CPU path (TF1 is not installable, so
  the op-for-op torch-CPU port of oracle/cpu_port.py) on the host cores."""
  rank = int(os.environ.get('RANK', '0'))
  if rank != 0:
    return
  T, A = args.T, args.A
  B = pick_cpu_sample(T, A, args.steps, args.warmup, budget_s=90.0)
  rate, sec_per_step, cores = time_cpu_port(T, B, A, args.steps, args.warmup)
  sample = ('T=%d B=%d A=%d per step (workload is B=%d per GPU), float32, '
            'torch-CPU port of vtrace.py + experiment.py:324-343 with autograd '
            'backward, %d threads, %s' % (T, B, A, args.B, cores, cpu_model()))
  line = {
      'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': UNIT,
      'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
      'ms_per_step': sec_per_step * 1e3, 'higher_is_better': True,
      'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
      'data': 'synthetic',
      'config': workload_config(
          T, args.B // args.gpus if args.scaling == 'strong' else args.B, A,
          args.gpus, args.scaling),
      'cpu_baseline': {'value': rate, 'unit': UNIT, 'cores': cores,
                       'kind': 'port', 'sample': sample},
      'e2e': {'value': rate, 'unit': UNIT, 'h2d_bytes_per_step': 0,
              'd2h_bytes_per_step': 0},
      'gpu_launches': 0,
  }
  print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------
class FusedCall(object):
  """Pre-allocated outputs + one ctypes call into the C ABI per step."""

  def __init__(self, inp, T, B, A, device):
    import torch
    from scalable_agent_b200 import _lib
    self.lib = _lib.load()
    self._lib_mod = _lib
    f32 = dict(dtype=torch.float32, device=device)
    self.out = {
        'vs': torch.empty(T, B, **f32), 'pg_advantages': torch.empty(T, B, **f32),
        'log_rhos': torch.empty(T, B, **f32),
        'behaviour_action_log_probs': torch.empty(T, B, **f32),
        'target_action_log_probs': torch.empty(T, B, **f32),
        'd_target_logits': torch.empty(T, B, A, **f32),
        'd_values': torch.empty(T, B, **f32),
        'losses': torch.empty(4, **f32)}
    self.ws_bytes = self.lib.vtr_fused_workspace_bytes(T, B, A)
    self.ws = torch.empty(self.ws_bytes // 8, dtype=torch.float64, device=device)
    p = _lib.FusedParams()
    p.T, p.B, p.A = T, B, A
    for k in FUSED_KEYS:
      setattr(p, k, inp[k].data_ptr())
    p.clip_rho_threshold = p.clip_pg_rho_threshold = 1.0   # vtrace.py:74
    p.baseline_cost, p.entropy_cost = BASELINE_COST, ENTROPY_COST
    for k, v in self.out.items():
      setattr(p, k, v.data_ptr())
    self.p = p
    self.inp = inp
    self.stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    self.describe()

  def describe(self):
    """What a call launches, for the parameters as they are now."""
    count = ctypes.c_int(0)
    text = ctypes.create_string_buffer(512)
    rc = self.lib.vtr_fused_describe(ctypes.byref(self.p), 1, ctypes.byref(count), text, 512)
    if rc != 0:
      self._lib_mod.check(rc, 'vtr_fused_describe')
    self.launches_per_call = count.value     # kernels of this library per call
    self.plan = text.value.decode().strip()
    return self.plan

  def __call__(self, stream=None):
    rc = self.lib.vtr_from_logits_fused_f32(
        ctypes.byref(self.p), self.ws.data_ptr(), self.ws_bytes,
        self.stream if stream is None else stream)
    if rc != 0:
      self._lib_mod.check(rc, 'vtr_from_logits_fused_f32')


def time_heads(T, B, A, call, device, steps):
  """The heads kernel at rows = T*B, K = 256 (the reference's core width):
  HBM-bound, 4K bytes read + 4(A+1) written per row."""
  import torch
  from scalable_agent_b200 import _lib, heads
  lib = _lib.load()
  Kc, rows = 256, T * B
  g = torch.Generator(device=device).manual_seed(99)
  core = torch.randn(rows, Kc, generator=g, device=device)
  packed = heads.pack_head_weights(
      torch.randn(Kc, A, generator=g, device=device) / 16, torch.zeros(A, device=device),
      torch.randn(Kc, generator=g, device=device) / 16, torch.zeros(1, device=device))
  logits = torch.empty(rows, A, device=device)
  baseline = torch.empty(rows, device=device)
  stream = ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)

  def head():
    rc = lib.vtr_policy_head_f32(core.data_ptr(), packed.data_ptr(), rows, Kc, A,
                                 logits.data_ptr(), baseline.data_ptr(), stream)
    if rc != 0:
      _lib.check(rc, 'vtr_policy_head_f32')

  def timed(fn, n):
    for _ in range(3):
      fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
      fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n

  n = max(5, min(steps, 30))
  head_ms = timed(head, n)
  saved = (call.p.target_policy_logits, call.p.values)
  call.p.target_policy_logits, call.p.values = logits.data_ptr(), baseline.data_ptr()

  def block():
    head()
    call()
  block_ms = timed(block, n)
  call.p.target_policy_logits, call.p.values = saved
  peak, _ = measured_peak_gbs()
  nbytes = rows * (4 * Kc + 4 * (A + 1)) + 4 * Kc * (A + 1)
  achieved = nbytes / (head_ms * 1e-3) / 1e9
  return {
      'workload': 'policy + baseline heads (experiment.py:199-203), [T*B=%d, %d] x [%d, %d], fp32-accurate '
                  '(3xTF32 tcgen05 kernel)' % (rows, Kc, Kc, A + 1),
      'ms_per_step': head_ms, 'rows_per_s': rows / (head_ms * 1e-3),
      'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                   'frac': achieved / peak, 'algorithmic_bytes_per_launch': nbytes},
      'block_ms_per_step': block_ms,
      'block': 'heads kernel + fused V-trace/loss pass, back to back on one stream',
      'block_timesteps_per_s': rows / (block_ms * 1e-3)}


def run_gpu(args):
  import torch
  import torch.distributed as dist

  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local_rank = int(os.environ.get('LOCAL_RANK', '0'))
  if not torch.cuda.is_available():
    raise SystemExit('bench.py: no CUDA device; the product path has no CPU '
                     'fallback (use --impl reference for the CPU baseline)')
  torch.cuda.set_device(local_rank)
  device = torch.device('cuda', local_rank)
  if world > 1:
    dist.init_process_group('nccl', device_id=device)
  T, B, A = args.T, args.B, args.A
  if args.scaling == 'strong':          # --B is then the GLOBAL batch
    if B % (4 * world):
      raise SystemExit('bench.py --scaling strong: B must be a multiple of 4 * ranks')
    B //= world
  K, W = args.steps, args.warmup
  # The per-SM speed table behind the tile sizes is a running mean over an
  # SM's first 20 tiles and an exponential mean (weight 0.05) after that: an
  # unreported pre-pass lets it settle whatever --warmup is (its length is in
  # the line, run.untimed_pre_pass_steps).  Measured with --steps 20 --warmup 5:
  # 30 steps before the timed region 144.2 us per step, 100 steps 143.4, 300
  # steps 143.1 (a learner runs for millions of steps; 100 steps are 14 ms).
  PRE = max(0, int(os.environ.get('VTR_BENCH_PRE', '100')) - W)

  # Inputs + outputs of one step must not fit the 126 MB L2, or steps would be
  # served from it: small per-GPU batches (strong scaling) rotate over several
  # independent sets of tensors.
  step_bytes = T * B * (12 * A + 40) + 4 * B
  n_sets = 1 if step_bytes > 300e6 else int(300e6 // step_bytes) + 1
  calls = [FusedCall(synthetic_on_device(T, B, A, 1000 * rank + 1 + 17 * k, device),
                     T, B, A, device) for k in range(n_sets)]
  call, inp = calls[0], calls[0].inp
  turn = [0]
  gathered = torch.empty(world, 4, dtype=torch.float32, device=device)
  # The only cross-GPU traffic of the path: the 4 loss scalars, summed in rank
  # order (bitwise reproducible).  Preferred: this library's own exchange over
  # NVLink peer memory; NCCL all-gather if symmetric memory is unavailable.
  exchange, exchange_kind = None, 'none'
  if world > 1:
    exchange_kind = 'nccl all_gather'
    if args.exchange in ('fused', 'p2p'):
      try:
        from scalable_agent_b200.distributed import LossExchange
        exchange = LossExchange(device)
        exchange_kind = (
            'p2p stores over NVLink and the wait for the peers, both inside the fused kernel'
            if args.exchange == 'fused' else
            'p2p stores over NVLink (vtr_exchange_*), 2 launches')
      except Exception as e:  # pylint: disable=broad-except
        exchange_kind = 'nccl all_gather (p2p exchange unavailable: %s)' % (
            str(e).splitlines()[0][:80])
    ok = torch.tensor([1 if exchange is not None else 0], device=device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)   # all ranks must agree
    if not int(ok.item()):
      exchange = None

  def exchange_losses(call):
    if exchange is not None:
      return exchange.all_sum(call.out['losses'])
    dist.all_gather_into_tensor(gathered, call.out['losses'])
    return gathered

  fused_exchange = exchange is not None and args.exchange == 'fused'
  global_losses = torch.zeros(4, dtype=torch.float32, device=device)

  def step():
    call = calls[turn[0] % n_sets]
    turn[0] += 1
    if fused_exchange:
      # every step publishes its scalars and hands out the sums of the step
      # before (already in this rank's buffer): no rank waits for another
      exchange.attach(call.p, global_losses, lag=args.exchange_lag)
      call()
      return
    call()
    if world > 1:
      exchange_losses(call)

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  for _ in range(PRE + W):
    step()
  barrier()

  sampler = ClockSampler(physical_gpu_index(local_rank))
  sampler.start()
  # The timed region is bracketed by ONE pair of CUDA events on the launching
  # stream; per-step events would put two extra stream operations between
  # consecutive kernels (measured: +6 us per step).
  t_begin = torch.cuda.Event(enable_timing=True)
  t_end = torch.cuda.Event(enable_timing=True)
  barrier()
  t_begin.record()
  for i in range(K):
    step()
  t_end.record()
  barrier()
  total_ms = t_begin.elapsed_time(t_end)
  # keep the GPU busy a little longer if the region was too short to sample
  if sampler.ok and len(sampler.samples) < 5:
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 0.3:
      for _ in range(20):
        call()
      torch.cuda.synchronize()
    clock_note = 'timed region + 0.3 s probe of the same kernel'
  else:
    clock_note = 'timed region'
  sampler.stop()

  if world > 1:
    t = torch.tensor([total_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
  kernel_mean_ms = total_ms / K      # max over ranks of the mean time per call
  value = T * B * world * K / (total_ms * 1e-3)

  # ---- are the cross-GPU sums right? ---------------------------------------
  # Two more steps A, B with the exchange of the timed loop; the sums it hands
  # out must equal, bit for bit, the rank-ordered float64 sum of the ranks'
  # local losses gathered with NCCL.
  exchange_check = None
  if world > 1:
    exchange_check = {'ok': None, 'ranks': world,
                      'reason': 'no peer-memory exchange in this run (%s)' % exchange_kind}
  if world > 1 and exchange is not None:
    def rank_ordered_sum(local):
      dist.all_gather_into_tensor(gathered, local)
      total = gathered[0].double()
      for r in range(1, world):          # rank 0, 1, ...: the order the kernel uses
        total = total + gathered[r].double()
      return total.float()

    pairs = []
    if fused_exchange:
      call_a, call_b = calls[0], calls[1 % n_sets]
      exchange.attach(call_a.p, global_losses, lag=args.exchange_lag)
      call_a()
      local_a = call_a.out['losses'].clone()
      sums_a = global_losses.clone() if args.exchange_lag == 0 else None
      exchange.attach(call_b.p, global_losses, lag=args.exchange_lag)
      call_b()
      local_b = call_b.out['losses'].clone()
      if args.exchange_lag == 0:
        pairs = [(sums_a, local_a), (global_losses.clone(), local_b)]
      else:
        pairs = [(global_losses.clone(), local_a), (exchange.collect(), local_b)]
    else:
      call_a = calls[0]
      call_a()
      pairs = [(exchange.all_sum(call_a.out['losses']).clone(), call_a.out['losses'].clone())]
    worst, ok = 0.0, True
    for got, local in pairs:
      want = rank_ordered_sum(local)
      finite = bool(torch.isfinite(got).all().item())
      ok = ok and finite and torch.equal(got, want)
      if finite:
        worst = max(worst, float((got.double() - want.double()).abs().max().item()))
    flag = torch.tensor([1 if ok else 0], device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)       # every rank must agree
    exchange_check = {'ok': bool(int(flag.item())), 'ranks': world,
                      'max_abs_diff': worst, 'steps_checked': len(pairs),
                      'against': 'nccl all_gather + rank-ordered float64 sum, bitwise'}
    barrier()

  # ---- the advertised Python API with device-resident inputs ---------------
  # experiment.compute_vtrace_loss(...) + backward(): what a user of the
  # drop-in gets per step when the learner's tensors already live on the GPU.
  from scalable_agent_b200 import experiment
  api_in = {k: v for k, v in inp.items()}
  api_in['target_policy_logits'] = inp['target_policy_logits'].detach().requires_grad_(True)
  api_in['values'] = inp['values'].detach().requires_grad_(True)

  def api_step():
    api_in['target_policy_logits'].grad = None
    api_in['values'].grad = None
    out = experiment.compute_vtrace_loss(
        baseline_cost=BASELINE_COST, entropy_cost=ENTROPY_COST, **api_in)
    out.total_loss.backward()

  for _ in range(max(3, W)):
    api_step()
  barrier()
  a_begin = torch.cuda.Event(enable_timing=True)
  a_end = torch.cuda.Event(enable_timing=True)
  a_begin.record()
  h0 = time.perf_counter()
  for _ in range(K):
    api_step()
  host_s = time.perf_counter() - h0
  a_end.record()
  barrier()
  api_ms = a_begin.elapsed_time(a_end) / K
  if world > 1:
    t = torch.tensor([api_ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    api_ms = float(t.item())
  api_device = {
      'value': T * B * world / (api_ms * 1e-3), 'unit': UNIT,
      'ms_per_step': api_ms, 'host_ms_per_step': host_s / K * 1e3,
      'vs_kernel': api_ms / (total_ms / K),
      'api': 'scalable_agent_b200.experiment.compute_vtrace_loss + backward, '
             'device-resident inputs, CUDA events'}

  # The same step through compute_vtrace_loss_and_grads (no autograd node: the
  # losses and both gradients straight from the one launch), here and at
  # BASELINE.json's config 2, where the host side is all there is.
  def timed(fn, steps):
    for _ in range(5):
      fn()
    torch.cuda.synchronize()
    b = torch.cuda.Event(enable_timing=True)
    e = torch.cuda.Event(enable_timing=True)
    b.record()
    t0 = time.perf_counter()
    for _ in range(steps):
      fn()
    host = time.perf_counter() - t0
    e.record()
    torch.cuda.synchronize()
    return b.elapsed_time(e) / steps, host / steps * 1e3

  def direct_step(batch):
    return lambda: experiment.compute_vtrace_loss_and_grads(
        baseline_cost=BASELINE_COST, entropy_cost=ENTROPY_COST, **batch)

  def autograd_step(batch):
    def step():
      batch['target_policy_logits'].grad = None
      batch['values'].grad = None
      experiment.compute_vtrace_loss(
          baseline_cost=BASELINE_COST, entropy_cost=ENTROPY_COST, return_aux=False, **batch).total_loss.backward()
    return step

  if world == 1:
    d_ms, d_host = timed(direct_step(api_in), K)
    api_device['direct'] = {
        'ms_per_step': d_ms, 'host_ms_per_step': d_host, 'vs_kernel': d_ms / (total_ms / K),
        'api': 'experiment.compute_vtrace_loss_and_grads (losses + gradients, no autograd node)'}
    small = synthetic_on_device(100, 32, 9, 2, device)
    small['target_policy_logits'] = small['target_policy_logits'].detach().requires_grad_(True)
    small['values'] = small['values'].detach().requires_grad_(True)
    a_ms, a_host = timed(autograd_step(small), 300)
    s_ms, s_host = timed(direct_step(small), 300)
    api_device['config2'] = {
        'workload': 'T=100 B=32 A=9', 'autograd_ms_per_step': a_ms, 'autograd_host_ms_per_step': a_host,
        'direct_ms_per_step': s_ms, 'direct_host_ms_per_step': s_host}
  del api_in

  # ---- the heads upstream of the path (experiment.py:199-203) ---------------
  # A second shape with a roofline of its own: [T*B, 256] x [256, A+1] on the
  # tcgen05 kernel, alone and in front of the fused pass (1 GPU only: it reads
  # 3.4 GB of core output at the headline shape).
  heads_line = None
  if world == 1 and not args.no_heads:
    try:
      heads_line = time_heads(T, B, A, calls[0], device, K)
    except Exception as e:  # pylint: disable=broad-except
      heads_line = {'error': str(e).splitlines()[0][:200]}

  # ---- end to end through the public API with host buffers ---------------
  e2e_steps = max(1, min(K, 10))
  host = {k: torch.empty(v.shape, dtype=v.dtype, pin_memory=True).copy_(v)
          for k, v in inp.items()}
  h2d = sum(v.numel() * v.element_size() for v in host.values())

  def e2e_step():
    dev_in = {k: v.to(device, non_blocking=True) for k, v in host.items()}
    dev_in['target_policy_logits'].requires_grad_(True)
    dev_in['values'].requires_grad_(True)
    out = experiment.compute_vtrace_loss(
        baseline_cost=BASELINE_COST, entropy_cost=ENTROPY_COST, **dev_in)
    out.total_loss.backward()
    result = torch.stack([out.total_loss.detach(), out.policy_gradient_loss,
                          out.baseline_loss, out.entropy_loss])
    if world > 1:
      dist.all_gather_into_tensor(gathered, result)
      result = gathered.sum(0)
    return result.cpu()      # device -> host read of the step's result

  for _ in range(2):
    e2e_step()
  barrier()
  t0 = time.perf_counter()
  for _ in range(e2e_steps):
    losses_host = e2e_step()
  barrier()
  e2e_s = time.perf_counter() - t0
  if world > 1:
    t = torch.tensor([e2e_s], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
  e2e_value = T * B * world * e2e_steps / e2e_s
  d2h = losses_host.numel() * losses_host.element_size()

  line = None
  if rank == 0:
    peak, peak_src = measured_peak_gbs()
    alg = algorithmic_bytes(T, B, A)
    achieved = alg / (kernel_mean_ms * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel is
    # not measurable from inside this process: it is copied from the last
    # `ncu --set full` capture of this shape (refreshed whenever K1c changes).
    traffic, traffic_src = None, None
    try:
      with open(os.path.join(ROOT, 'profiles', 'dram_traffic.json')) as f:
        rec = json.load(f)
      if rec.get('T') == T and rec.get('B') == B and rec.get('A') == A:
        traffic = rec.get('dram_bytes_per_launch')
        traffic_src = 'profiles/dram_traffic.json <- %s' % rec.get('source', 'ncu capture')
    except Exception:  # pylint: disable=broad-except
      pass
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world,
        'steps': K, 'warmup': W, 'ms_per_step': total_ms / K,
        'higher_is_better': True, 'scaling': args.scaling, 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': workload_config(T, B, A, world, args.scaling),
        'run': {
            'parallelism': ('single GPU' if world == 1 else
                            'B sharded over %d ranks (%d columns each); sum of 4 '
                            'loss scalars per step via %s' % (world, B, exchange_kind)),
            'l2': ('no flush: %.0f MB of inputs + %.0f MB of outputs per step '
                   'exceed the 126 MB L2' % (
                       (T * B * (8 * A + 16) + 4 * B) / 1e6,
                       T * B * (4 * A + 24) / 1e6) if n_sets == 1 else
                   'no flush: steps rotate over %d independent sets of tensors, '
                   '%.0f MB in all, more than the 126 MB L2' % (
                       n_sets, n_sets * step_bytes / 1e6)),
            'untimed_pre_pass_steps': PRE},
        'roofline': {
            'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
            'frac': achieved / peak, 'traffic': traffic,
            'traffic_source': traffic_src,
            'peak_source': peak_src,
            'algorithmic_bytes_per_launch': alg,
            'kernel': call.plan + ' (time = timed region / steps: the whole '
                      'call, incl. the cursor memset%s)' % (
                          ' and the loss exchange' if world > 1 else ''),
            'kernel_ms_mean': kernel_mean_ms,
            'frac_of_nominal_8TBs': achieved / 8000.0},
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d,
                'd2h_bytes_per_step': d2h, 'steps': e2e_steps,
                'ms_per_step': e2e_s / e2e_steps * 1e3,
                'api': 'scalable_agent_b200.experiment.compute_vtrace_loss + '
                       'backward, pinned host inputs'},
        'api_device': api_device,
        'exchange_check': exchange_check,
        'heads': heads_line,
        'gpu_launches': (call.launches_per_call +
                         (0 if exchange is None or fused_exchange else 2)) * K,
        'clocks': dict(sampler.summary(), sampled=clock_note),
    }
  if world > 1:
    dist.barrier()
  if rank == 0:
    if world == 1 and not args.no_cpu_baseline:
      Bc = pick_cpu_sample(T, A, 6, 1, budget_s=20.0)
      rate, _, cores = time_cpu_port(T, Bc, A, 6, 1)
      line['cpu_baseline'] = {
          'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port',
          'sample': '6 steps of T=%d B=%d A=%d, float32 torch-CPU port of the '
                    'reference path with autograd backward (oracle/cpu_port.py)'
                    ', %s' % (T, Bc, A, cpu_model())}
    else:
      line['cpu_baseline'] = None
    print(json.dumps(line), flush=True)
  if world > 1:
    dist.destroy_process_group()
  if exchange_check is not None and exchange_check.get('ok') is False:
    raise SystemExit('bench.py: the cross-GPU loss sums do not match the NCCL '
                     'reference: %s' % json.dumps(exchange_check))


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--gpus', type=int, default=1)
  ap.add_argument('--steps', type=int, default=200)
  ap.add_argument('--warmup', type=int, default=20)
  ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
  ap.add_argument('--T', type=int, default=T_DEFAULT)
  ap.add_argument('--B', type=int, default=B_DEFAULT, help='columns per GPU')
  ap.add_argument('--A', type=int, default=A_DEFAULT)
  ap.add_argument('--scaling', default='weak', choices=['weak', 'strong'],
                  help='weak: --B columns on every GPU; strong: --B columns in all')
  ap.add_argument('--no-cpu-baseline', action='store_true')
  ap.add_argument('--no-heads', action='store_true',
                  help='skip the heads-kernel leg (tcgen05 contraction upstream of the path)')
  ap.add_argument('--exchange', default='fused', choices=['fused', 'p2p', 'nccl'],
                  help='how the loss scalars are summed over ranks (N > 1)')
  ap.add_argument('--exchange-lag', type=int, default=1, choices=[0, 1],
                  help='fused exchange: 1 = each step hands out the sums of the '
                       'step before (no rendezvous inside the step), 0 = its own')
  args = ap.parse_args()
  if args.impl == 'reference':
    run_reference(args)
  else:
    run_gpu(args)


if __name__ == '__main__':
  main()
