import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

GOLDEN_PATH = os.path.join(ROOT, 'tests', 'golden', 'vtrace_golden.npz')


def pytest_configure(config):
  config.addinivalue_line(
      'markers', 'gpu: needs a CUDA device (run on the B200 box via gpurun)')
  # The C-ABI library is a build artefact (git-ignored).  Build it once if a
  # fresh checkout has none: nvcc cross-compiles sm_100a without a GPU.
  lib = os.path.join(ROOT, 'scalable_agent_b200', 'lib', 'libvtrace_b200.so')
  if not os.path.exists(lib) and not os.environ.get('VTRACE_B200_LIB'):
    import subprocess
    subprocess.run(['make', '-C', os.path.join(ROOT, 'scalable_agent_b200', 'csrc'),
                    '-j', str(max(1, min(8, os.cpu_count() or 1)))], check=True)


class Golden(object):
  """tests/golden/vtrace_golden.npz, grouped by case name."""

  def __init__(self, path=GOLDEN_PATH):
    self._npz = np.load(path)
    self.cases = {}
    for key in self._npz.files:
      case, field = key.split('/', 1)
      self.cases.setdefault(case, {})[field] = self._npz[key]

  def case(self, name):
    return dict(self.cases[name])

  @staticmethod
  def threshold(value):
    """NaN encodes None (no clipping) in the fixture."""
    value = float(value)
    return None if np.isnan(value) else value


@pytest.fixture(scope='session')
def golden():
  return Golden()


def assert_close(actual, expected, tol=1e-5, what=''):
  """|x - ref| <= tol * max(1, |ref|) elementwise."""
  actual = np.asarray(actual, dtype=np.float64)
  expected = np.asarray(expected, dtype=np.float64)
  assert actual.shape == expected.shape, (what, actual.shape, expected.shape)
  err = np.abs(actual - expected) / np.maximum(1.0, np.abs(expected))
  worst = float(err.max()) if err.size else 0.0
  assert np.all(np.isfinite(actual)), '%s: non-finite values' % what
  assert worst <= tol, '%s: max scaled error %.3e > %.1e' % (what, worst, tol)
