"""Seeded random sweep of the fused pass: shapes, kernels, thresholds, reward
clipping, done flags and loss weights drawn at random (fixed seeds, so a
failure reproduces), every output against the float64 oracle at 1e-5.

The hand-picked cases of test_vtrace_gpu.py sit on the edges the launch
heuristics have today; this file is there for the edges they will have
tomorrow (a short last tile, a batch one column past a tile, more time steps
than ring slots, an unaligned row, ...).

Tolerance: 1e-5 scaled, the bound of every other GPU test, whenever both
importance weights are clipped at 1 or below (the reference's defaults,
vtrace.py:74).  With a looser or no clip the weights reach 8-14 on these
inputs, v_s grows with them and float32 itself gives out earlier: the
op-for-op float32 port of the reference (oracle/cpu_port.py) is 1.1e-5 ...
2.1e-5 away from the float64 oracle on exactly these configurations.  That is
asserted, not assumed: for every such case the port is run on the same inputs
and the kernel is held to max(1e-5, 2 x the port's own error), never more than
5e-5.  (Both are float32 evaluations of the same expression with different
rounding sequences - torch's vectorised expf and one pass per op there, ex2 /
FMA and one fused pass here - so neither is uniformly the closer one; the
largest kernel / port ratio seen over the 48 cases is 1.50, at 1.7e-5.)"""
import numpy as np
import pytest
import torch

from oracle import cpu_port
from oracle import vtrace_oracle as oracle
from tests.test_vtrace_gpu import FUSED_KEYS, _check_fused

pytestmark = pytest.mark.gpu

SPECIAL_A = (3, 4, 6, 9, 18)


def _scaled_err(x, ref):
  x = np.asarray(x, dtype=np.float64)
  ref = np.asarray(ref, dtype=np.float64)
  return float(np.max(np.abs(x - ref) / np.maximum(1.0, np.abs(ref)))) if ref.size else 0.0


def _port_error(inp, crho, cpg, bc, ec, clip, done, discounting):
  """Largest scaled error of the float32 op-for-op port of the reference
  against the float64 oracle, over every output the kernel is checked on."""
  rewards = oracle.clip_rewards(inp['rewards'], clip).astype(np.float32)
  discounts = inp['discounts']
  if done is not None:
    discounts = oracle.discounts_from_done(done, discounting).astype(np.float32)
  args = dict(inp, rewards=rewards, discounts=discounts)
  ref = oracle.fused_vtrace_loss(*[args[k] for k in FUSED_KEYS], crho, cpg, bc, ec)
  torch.set_num_threads(max(1, torch.get_num_threads()))
  losses, d_logits, d_values, vs, pg = cpu_port.learner_loss(
      *[torch.from_numpy(np.ascontiguousarray(args[k])) for k in FUSED_KEYS],
      baseline_cost=bc, entropy_cost=ec, clip_rho_threshold=crho,
      clip_pg_rho_threshold=cpg)
  total = float(losses[0]) + bc * float(losses[1]) + ec * float(losses[2])
  return max(
      _scaled_err(vs.numpy(), ref.vs), _scaled_err(pg.numpy(), ref.pg_advantages),
      _scaled_err(d_logits.numpy(), ref.d_target_logits),
      _scaled_err(d_values.numpy(), ref.d_values),
      _scaled_err(losses.numpy(), [ref.pg_loss, ref.baseline_loss, ref.entropy_loss]),
      _scaled_err(total, ref.total_loss))


def _draw(rng):
  kind = rng.integers(0, 4)
  if kind == 0:      # tiny
    T, B = int(rng.integers(1, 8)), int(rng.integers(1, 40))
  elif kind == 1:    # a few tiles, ragged
    T, B = int(rng.integers(2, 70)), int(rng.integers(33, 400))
  elif kind == 2:    # around the per-SM kernel's threshold (64 columns per SM)
    T, B = int(rng.integers(30, 45)), int(148 * 64 + 4 * rng.integers(-3, 40))
  else:              # long unroll, small batch
    T, B = int(rng.integers(100, 257)), int(rng.integers(1, 70))
  A = int(rng.choice(SPECIAL_A)) if rng.random() < 0.6 else int(rng.integers(1, 41))
  return T, B, A


@pytest.mark.parametrize('seed', range(48))
def test_random_configuration_matches_oracle(seed):
  assert torch.cuda.is_available()
  from scalable_agent_b200 import experiment
  rng = np.random.default_rng(1000 + seed)
  T, B, A = _draw(rng)
  paths = ['auto', 'column']
  if B % 4 == 0:
    paths.append('sm_tile')
  if T <= 256 and A in SPECIAL_A and B <= 16384:
    paths.append('item')
  if T <= 100 and B <= 16384 and A <= 40:   # chunks of <= 13 steps, a portable cluster
    paths.append('chunk')
  path = paths[int(rng.integers(0, len(paths)))]
  crho = [None, 1.0, 3.7, 0.5][int(rng.integers(0, 4))]
  cpg = [None, 1.0, 2.2][int(rng.integers(0, 3))]
  bc = float(rng.choice([0.5, 1.0, 0.0]))
  ec = float(rng.choice([0.00025, 0.0033391318945337044, 0.0, 0.01]))
  clip = [None, 'abs_one', 'soft_asymmetric'][int(rng.integers(0, 3))]
  use_done = bool(rng.integers(0, 2))
  inp = oracle.synthetic_inputs(T, B, A, seed=seed, reward_clipping=None)
  inp['rewards'] = inp['raw_rewards']
  discounting = float(rng.choice([0.99, 0.9, 1.0])) if use_done else 0.99
  tol = 1e-5
  if not ((crho == 1.0 or crho == 0.5) and cpg == 1.0):
    port_err = _port_error(inp, crho, cpg, bc, ec, clip,
                           inp['done'] if use_done else None, discounting)
    tol = max(1e-5, 2.0 * port_err)
    assert tol <= 5e-5, 'the float32 port itself is %.2e off on this case' % port_err
  kw = dict(crho=crho, cpg=cpg, bc=bc, ec=ec, path=path, tol=tol)
  if clip is not None:
    kw['reward_clipping'] = clip
  if use_done:
    kw['done_np'] = inp['done']
    kw['discounting'] = discounting
  try:
    _check_fused(experiment, inp, **kw)
  except AssertionError as e:
    raise AssertionError('T=%d B=%d A=%d path=%s crho=%s cpg=%s bc=%s ec=%s clip=%s done=%s: %s'
                         % (T, B, A, path, crho, cpg, bc, ec, clip, use_done, e))
