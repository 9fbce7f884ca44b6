"""Freeze the reference's own expected values as fixtures.

Run in the build container (where /root/reference is mounted):

    python tests/golden/make_golden.py

Every EXPECTED array written here comes out of code lifted from
/root/reference/vtrace_test.py by oracle/lift_reference.py (`_shaped_arange`,
`_softmax`, `_ground_truth_calculation`); nothing from this repository's own
oracle or kernels is on the expected side.  The INPUTS follow the recipes of
the reference tests (cited per case) with the unseeded `np.random.randint`
actions replaced by seeded ones, plus a few larger synthetic batches.

Output: tests/golden/vtrace_golden.npz  (keys "<case>/<field>").
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import lift_reference  # noqa: E402
from oracle import vtrace_oracle  # noqa: E402  (input generator only)


def _gather(arr, actions):
  return np.take_along_axis(arr, actions[..., None].astype(np.int64), -1)[..., 0]


def main():
  ref = lift_reference.lift()
  out = {}

  # --- vtrace_test.py:120-156  test_vtrace, Batch1 / Batch5 -----------------
  for batch_size in (1, 5):
    seq_len = 5
    log_rhos = ref._shaped_arange(seq_len, batch_size) / (batch_size * seq_len)
    log_rhos = 5 * (log_rhos - 0.5)
    values = {
        'log_rhos': log_rhos,
        'discounts': np.array([[0.9 / (b + 1) for b in range(batch_size)]
                               for _ in range(seq_len)]),
        'rewards': ref._shaped_arange(seq_len, batch_size),
        'values': ref._shaped_arange(seq_len, batch_size) / batch_size,
        'bootstrap_value': ref._shaped_arange(batch_size) + 1.0,
        'clip_rho_threshold': 3.7,
        'clip_pg_rho_threshold': 2.2,
    }
    gt = ref._ground_truth_calculation(**values)          # as the test runs it
    values64 = {k: (np.asarray(v, np.float64) if isinstance(v, np.ndarray)
                    else v) for k, v in values.items()}
    gt64 = ref._ground_truth_calculation(**values64)      # float64 inputs
    case = 'test_vtrace_B%d' % batch_size
    for k, v in values.items():
      out['%s/%s' % (case, k)] = np.asarray(v)
    out[case + '/vs'] = gt.vs
    out[case + '/pg_advantages'] = gt.pg_advantages
    out[case + '/vs_f64'] = gt64.vs
    out[case + '/pg_advantages_f64'] = gt64.pg_advantages

  # --- vtrace_test.py:89-115  test_log_probs_from_logits_and_actions --------
  rng = np.random.RandomState(7)
  for batch_size in (1, 2):
    seq_len, num_actions = 7, 3
    policy_logits = ref._shaped_arange(seq_len, batch_size, num_actions) + 10
    actions = rng.randint(0, num_actions, size=(seq_len, batch_size),
                          dtype=np.int32)
    mask = actions[..., None] == np.arange(num_actions)
    expected = np.log(ref._softmax(policy_logits))[mask].reshape(
        seq_len, batch_size)
    case = 'test_log_probs_B%d' % batch_size
    out[case + '/policy_logits'] = policy_logits
    out[case + '/actions'] = actions
    out[case + '/log_probs'] = expected

  # --- vtrace_test.py:158-247  test_vtrace_from_logits (config 1) -----------
  for batch_size in (1, 2):
    seq_len, num_actions = 5, 3
    logits = ref._shaped_arange(seq_len, batch_size, num_actions)
    actions = rng.randint(0, num_actions - 1,
                          size=(seq_len, batch_size)).astype(np.int32)
    discounts = np.array([[0.9 / (b + 1) for b in range(batch_size)]
                          for _ in range(seq_len)])
    rewards = ref._shaped_arange(seq_len, batch_size)
    values = ref._shaped_arange(seq_len, batch_size) / batch_size
    bootstrap = ref._shaped_arange(batch_size) + 1.0
    lp = _gather(np.log(ref._softmax(logits.astype(np.float64))), actions)
    log_rhos = lp - lp
    gt = ref._ground_truth_calculation(
        discounts=discounts.astype(np.float64), log_rhos=log_rhos,
        rewards=rewards.astype(np.float64), values=values.astype(np.float64),
        bootstrap_value=bootstrap.astype(np.float64),
        clip_rho_threshold=None, clip_pg_rho_threshold=None)
    case = 'test_vtrace_from_logits_B%d' % batch_size
    out[case + '/behaviour_policy_logits'] = logits
    out[case + '/target_policy_logits'] = logits
    out[case + '/actions'] = actions
    out[case + '/discounts'] = discounts
    out[case + '/rewards'] = rewards
    out[case + '/values'] = values
    out[case + '/bootstrap_value'] = bootstrap
    out[case + '/vs'] = gt.vs
    out[case + '/pg_advantages'] = gt.pg_advantages
    out[case + '/log_rhos'] = log_rhos
    out[case + '/behaviour_action_log_probs'] = lp
    out[case + '/target_action_log_probs'] = lp

  # --- larger synthetic batches (SURVEY.md 8(d) recipe), rho != 1 -----------
  # (name, T, B, A, clip_rho, clip_pg_rho, reward clipping, seed)
  synthetic = [
      ('synthetic_T100_B32_A9', 100, 32, 9, 1.0, 1.0, 'abs_one', 1),
      ('synthetic_T20_B96_A18', 20, 96, 18, 1.0, 1.0, 'abs_one', 2),
      ('synthetic_T100_B40_A18_noclip', 100, 40, 18, None, None,
       'soft_asymmetric', 3),
      ('synthetic_T33_B7_A5_clip', 33, 7, 5, 3.7, 2.2, 'abs_one', 4),
  ]
  for name, T, B, A, crho, cpg, rclip, seed in synthetic:
    inp = vtrace_oracle.synthetic_inputs(T, B, A, seed=seed,
                                         reward_clipping=rclip)
    b_lp = _gather(np.log(ref._softmax(
        inp['behaviour_policy_logits'].astype(np.float64))), inp['actions'])
    t_lp = _gather(np.log(ref._softmax(
        inp['target_policy_logits'].astype(np.float64))), inp['actions'])
    log_rhos = t_lp - b_lp
    gt = ref._ground_truth_calculation(
        discounts=inp['discounts'].astype(np.float64), log_rhos=log_rhos,
        rewards=inp['rewards'].astype(np.float64),
        values=inp['values'].astype(np.float64),
        bootstrap_value=inp['bootstrap_value'].astype(np.float64),
        clip_rho_threshold=crho, clip_pg_rho_threshold=cpg)
    for k in ('behaviour_policy_logits', 'target_policy_logits', 'actions',
              'discounts', 'rewards', 'values', 'bootstrap_value'):
      out['%s/%s' % (name, k)] = inp[k]
    out[name + '/clip_rho_threshold'] = np.float64(
        np.nan if crho is None else crho)      # NaN encodes None
    out[name + '/clip_pg_rho_threshold'] = np.float64(
        np.nan if cpg is None else cpg)
    out[name + '/vs'] = gt.vs
    out[name + '/pg_advantages'] = gt.pg_advantages
    out[name + '/log_rhos'] = log_rhos
    out[name + '/behaviour_action_log_probs'] = b_lp
    out[name + '/target_action_log_probs'] = t_lp

  path = os.path.join(HERE, 'vtrace_golden.npz')
  np.savez_compressed(path, **out)
  print('wrote %s: %d arrays, %d bytes' % (path, len(out),
                                           os.path.getsize(path)))
  print('anchors  B1 vs[0]=%r pg[0]=%r' % (
      out['test_vtrace_B1/vs'][0, 0], out['test_vtrace_B1/pg_advantages'][0, 0]))
  print('anchors  B5 vs[0,:3]=%r pg[0,:3]=%r' % (
      out['test_vtrace_B5/vs'][0, :3], out['test_vtrace_B5/pg_advantages'][0, :3]))


if __name__ == '__main__':
  main()
