"""Size-independent properties of the path at the full benchmark shape
(T=100, B=32768, A=18), where the float64 oracle is too slow to be the only
check.  Each property follows from the reference's formulas and is verified
on the kernel's own outputs with torch arithmetic on the device:

* on-policy (behaviour logits == target logits, clips >= 1): log_rhos == 0,
  v_s obeys v_s(t) = r_t + gamma_t v_s(t+1) (the n-step Bellman recursion
  V-trace reduces to, vtrace.py:235-265) and pg_advantages == v_s - V;
* every row of d(total)/d(logits) sums to zero (a softmax gradient);
* the loss scalars equal the sums of experiment.py:328,335,343 recomputed in
  float64 from the kernel's per-item outputs (a checksum of checksums);
* adding a constant to every logit of a row changes nothing (softmax shift
  invariance);
* d(total)/d(values) == baseline_cost * (V - v_s) (experiment.py:404-405).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

T, B, A = 100, 32768, 18
BC, EC = 0.5, 0.00025


@pytest.fixture(scope='module')
def batch():
  assert torch.cuda.is_available()
  g = torch.Generator(device='cuda').manual_seed(7)
  f = dict(device='cuda', dtype=torch.float32, generator=g)
  behaviour = torch.randn(T, B, A, **f)
  target = behaviour + 0.5 * torch.randn(T, B, A, **f)
  actions = torch.randint(0, A, (T, B), device='cuda', generator=g, dtype=torch.int32)
  done = torch.rand(T, B, device='cuda', generator=g) < 0.01
  discounts = (~done).float() * 0.99
  rewards = (2 * torch.randn(T, B, **f)).clamp(-1, 1)
  values = torch.randn(T, B, **f)
  bootstrap = torch.randn(B, **f)
  return dict(behaviour=behaviour, target=target, actions=actions, discounts=discounts,
              rewards=rewards, values=values, bootstrap=bootstrap)


def _run(batch, behaviour=None, target=None):
  from scalable_agent_b200 import experiment
  target = (batch['target'] if target is None else target).detach().clone().requires_grad_(True)
  values = batch['values'].detach().clone().requires_grad_(True)
  behaviour = batch['behaviour'] if behaviour is None else behaviour
  out = experiment.compute_vtrace_loss(
      behaviour, target, batch['actions'], batch['discounts'], batch['rewards'], values,
      batch['bootstrap'], baseline_cost=BC, entropy_cost=EC, return_aux=True)
  out.total_loss.backward()
  return out, target.grad, values.grad


def test_on_policy_reduces_to_n_step_returns(batch):
  out, _, _ = _run(batch, behaviour=batch['target'])
  vt = out.vtrace_returns
  assert float(vt.log_rhos.abs().max()) == 0.0
  vs = vt.vs.double()
  vs_next = torch.cat([vs[1:], batch['bootstrap'].double()[None]], 0)
  bellman = batch['rewards'].double() + batch['discounts'].double() * vs_next
  err = ((vs - bellman).abs() / bellman.abs().clamp(min=1)).max()
  assert float(err) <= 1e-5, float(err)
  adv = vs - batch['values'].double()
  err = ((vt.pg_advantages.double() - adv).abs() / adv.abs().clamp(min=1)).max()
  assert float(err) <= 1e-5, float(err)


def test_gradient_rows_sum_to_zero_and_value_gradient(batch):
  out, d_logits, d_values = _run(batch)
  row_sums = d_logits.double().sum(-1).abs()
  scale = d_logits.double().abs().sum(-1).clamp(min=1)
  assert float((row_sums / scale).max()) <= 1e-6
  want = BC * (batch['values'].double() - out.vtrace_returns.vs.double())
  err = ((d_values.double() - want).abs() / want.abs().clamp(min=1)).max()
  assert float(err) <= 1e-6, float(err)


def test_loss_scalars_are_the_sums_of_the_per_item_outputs(batch):
  out, _, _ = _run(batch)
  vt = out.vtrace_returns
  pg = -(vt.target_action_log_probs.double() * vt.pg_advantages.double()).sum()      # :338-343
  bl = 0.5 * ((vt.vs.double() - batch['values'].double()) ** 2).sum()               # :324-328
  logp = torch.log_softmax(batch['target'].double(), -1)
  ent = (logp.exp() * logp).sum()                                                   # :331-335
  for got, want, what in [(out.policy_gradient_loss, pg, 'pg'), (out.baseline_loss, bl, 'baseline'),
                          (out.entropy_loss, ent, 'entropy'),
                          (out.total_loss, pg + BC * bl + EC * ent, 'total')]:
    got, want = float(got.detach()), float(want)
    assert abs(got - want) / max(1.0, abs(want)) <= 1e-6, (what, got, want)


def test_softmax_shift_invariance(batch):
  out, d_logits, d_values = _run(batch)
  shift = torch.randn(T, B, 1, device='cuda', generator=torch.Generator(device='cuda').manual_seed(3))
  out2, d_logits2, d_values2 = _run(batch, behaviour=batch['behaviour'] - 2 * shift,
                                    target=batch['target'] + shift)
  def close(a, b, tol, what):
    a, b = a.detach().double(), b.detach().double()
    err = float(((a - b).abs() / b.abs().clamp(min=1)).max())
    assert err <= tol, (what, err)
  close(out2.vtrace_returns.vs, out.vtrace_returns.vs, 2e-5, 'vs')
  close(out2.vtrace_returns.pg_advantages, out.vtrace_returns.pg_advantages, 2e-5, 'pg')
  close(d_logits2, d_logits, 2e-5, 'd_logits')
  close(d_values2, d_values, 2e-5, 'd_values')
  close(out2.total_loss, out.total_loss, 1e-5, 'total')
