"""Lift the reference's own numpy ground truth out of its test file.  TEST INFRASTRUCTURE ONLY.

``/root/reference/vtrace_test.py`` cannot be imported here (its top level does
``import tensorflow`` and ``import vtrace``, and TensorFlow 1.x is not
installable in this image), but the three helpers it defines for its expected
values are plain numpy:

* ``_shaped_arange``            vtrace_test.py:33-35
* ``_softmax``                  vtrace_test.py:38-40
* ``_ground_truth_calculation`` vtrace_test.py:44-83

This module parses the file, compiles only those three function definitions
into a scratch namespace (nothing is copied into this repository) and hands
them back.  It only works where the reference tree is mounted, i.e. in the
build container; the GPU box has no ``/root/reference``, which is why
``tests/golden/make_golden.py`` freezes the outputs as fixtures.
"""
import ast
import collections
import os
import types

import numpy as np

REFERENCE_TEST = '/root/reference/vtrace_test.py'
_WANTED = ('_shaped_arange', '_softmax', '_ground_truth_calculation')


def available():
  return os.path.exists(REFERENCE_TEST)


def lift(path=REFERENCE_TEST):
  """Returns a namespace with the reference's three numpy helpers."""
  with open(path, 'r') as f:
    tree = ast.parse(f.read(), filename=path)
  picked = [node for node in tree.body
            if isinstance(node, ast.FunctionDef) and node.name in _WANTED]
  missing = set(_WANTED) - {n.name for n in picked}
  if missing:
    raise RuntimeError('reference test file lacks %s' % sorted(missing))
  module = ast.Module(body=picked, type_ignores=[])
  # The ground truth returns ``vtrace.VTraceReturns``; give it a stand-in
  # module holding just that namedtuple (vtrace.py:42).
  vtrace_stub = types.SimpleNamespace(
      VTraceReturns=collections.namedtuple('VTraceReturns', 'vs pg_advantages'))
  namespace = {'np': np, 'vtrace': vtrace_stub}
  exec(compile(module, path, 'exec'), namespace)  # pylint: disable=exec-used
  return types.SimpleNamespace(**{k: namespace[k] for k in _WANTED})
