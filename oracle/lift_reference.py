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


# The reference tree is untrusted content and the lifted functions are
# executed, so their syntax trees are checked first: plain arithmetic over
# numpy, nothing else.  Allowed node types, callables and attribute roots:
_ALLOWED_NODES = (
    ast.FunctionDef, ast.arguments, ast.arg, ast.Return, ast.Assign, ast.AugAssign,
    ast.Expr, ast.For, ast.If, ast.IfExp, ast.Pass, ast.Name, ast.Load, ast.Store,
    ast.Constant, ast.Tuple, ast.List, ast.Subscript, ast.Slice, ast.Index if hasattr(ast, 'Index') else ast.Slice,
    ast.BinOp, ast.UnaryOp, ast.BoolOp, ast.Compare, ast.Call, ast.Attribute, ast.keyword,
    ast.Add, ast.Sub, ast.Mult, ast.Div, ast.FloorDiv, ast.Mod, ast.Pow, ast.USub, ast.UAdd,
    ast.Not, ast.And, ast.Or, ast.Eq, ast.NotEq, ast.Lt, ast.LtE, ast.Gt, ast.GtE, ast.Is,
    ast.IsNot, ast.In, ast.NotIn, ast.Starred)
_ALLOWED_CALL_NAMES = frozenset(_WANTED) | {'range', 'reversed', 'len', 'float', 'int', 'zip', 'enumerate'}
_ALLOWED_ATTR_ROOTS = frozenset({'np', 'vtrace'})
_FORBIDDEN_NAMES = frozenset({
    'exec', 'eval', 'open', 'compile', '__import__', 'getattr', 'setattr', 'delattr', 'globals',
    'locals', 'vars', 'input', 'breakpoint', 'exit', 'quit', 'help', 'memoryview', 'type', 'object'})


def _check_tree(node):
  """Raises unless `node` (a lifted function) only does arithmetic on numpy."""
  for child in ast.walk(node):
    if not isinstance(child, _ALLOWED_NODES):
      raise RuntimeError('lifted reference code contains a %s node' % type(child).__name__)
    if isinstance(child, ast.Name) and (child.id in _FORBIDDEN_NAMES or child.id.startswith('__')):
      raise RuntimeError('lifted reference code names %r' % child.id)
    if isinstance(child, ast.Attribute):
      if child.attr.startswith('_'):
        raise RuntimeError('lifted reference code reaches for attribute %r' % child.attr)
      root = child
      while isinstance(root, ast.Attribute):
        root = root.value
      # attributes of np / vtrace, or methods of a local array (x.reshape, x.shape ...)
      if isinstance(root, ast.Name) and root.id in _FORBIDDEN_NAMES:
        raise RuntimeError('lifted reference code reaches into %r' % root.id)
    if isinstance(child, ast.Call) and isinstance(child.func, ast.Name) \
        and child.func.id not in _ALLOWED_CALL_NAMES:
      raise RuntimeError('lifted reference code calls %r' % child.func.id)
    if isinstance(child, ast.Call) and isinstance(child.func, ast.Attribute):
      root = child.func
      while isinstance(root, ast.Attribute):
        root = root.value
      if isinstance(root, ast.Name) and root.id == 'np' and child.func.attr in (
          'load', 'save', 'savez', 'fromfile', 'tofile', 'memmap', 'loadtxt', 'savetxt', 'genfromtxt',
          'frombuffer', 'ctypeslib', 'lib', 'DataSource'):
        raise RuntimeError('lifted reference code calls np.%s' % child.func.attr)


def lift(path=REFERENCE_TEST):
  """Returns a namespace with the reference's three numpy helpers."""
  with open(path, 'r') as f:
    tree = ast.parse(f.read(), filename=path)
  picked = [node for node in tree.body
            if isinstance(node, ast.FunctionDef) and node.name in _WANTED]
  missing = set(_WANTED) - {n.name for n in picked}
  if missing:
    raise RuntimeError('reference test file lacks %s' % sorted(missing))
  for node in picked:
    if node.decorator_list:
      raise RuntimeError('lifted reference function %s is decorated' % node.name)
    _check_tree(node)
  module = ast.Module(body=picked, type_ignores=[])
  # The ground truth returns ``vtrace.VTraceReturns``; give it a stand-in
  # module holding just that namedtuple (vtrace.py:42).
  vtrace_stub = types.SimpleNamespace(
      VTraceReturns=collections.namedtuple('VTraceReturns', 'vs pg_advantages'))
  builtins_allowed = {k: getattr(__import__('builtins'), k) for k in
                      ('range', 'reversed', 'len', 'float', 'int', 'zip', 'enumerate')}
  namespace = {'np': np, 'vtrace': vtrace_stub, '__builtins__': builtins_allowed}
  exec(compile(module, path, 'exec'), namespace)  # pylint: disable=exec-used
  return types.SimpleNamespace(**{k: namespace[k] for k in _WANTED})
