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
them back.  ``lift_tf_path`` (further down) does the same for the TensorFlow
side of the path - ``vtrace.py`` and the loss block of ``experiment.py`` -
which it executes over ``oracle/tf_shim.py``.  It only works where the reference tree is mounted, i.e. in the
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


def _check_tree(node, extra_nodes=(), extra_calls=frozenset(), extra_attrs=frozenset()):
  """Raises unless `node` (a lifted function) only does arithmetic on numpy."""
  for child in ast.walk(node):
    if not isinstance(child, _ALLOWED_NODES + tuple(extra_nodes)):
      raise RuntimeError('lifted reference code contains a %s node' % type(child).__name__)
    if isinstance(child, ast.Name) and (child.id in _FORBIDDEN_NAMES or child.id.startswith('__')):
      raise RuntimeError('lifted reference code names %r' % child.id)
    if isinstance(child, ast.Attribute):
      if child.attr.startswith('_') and child.attr not in extra_attrs:
        raise RuntimeError('lifted reference code reaches for attribute %r' % child.attr)
      root = child
      while isinstance(root, ast.Attribute):
        root = root.value
      # attributes of np / vtrace, or methods of a local array (x.reshape, x.shape ...)
      if isinstance(root, ast.Name) and root.id in _FORBIDDEN_NAMES:
        raise RuntimeError('lifted reference code reaches into %r' % root.id)
    if isinstance(child, ast.Call) and isinstance(child.func, ast.Name) \
        and child.func.id not in _ALLOWED_CALL_NAMES | extra_calls:
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


# ---- the TensorFlow side of the path, run over oracle/tf_shim.py -------------
#
# vtrace.py (the module body: two namedtuples and the three public functions)
# and, from experiment.py, the three loss helpers (:324-343) plus the
# statements of build_learner that clip the rewards, form the discounts, call
# V-trace and weigh the total loss (:365-407).  The statements are lifted as
# they stand and wrapped into one function whose arguments are the names they
# read; nothing is copied into this repository.
REFERENCE_VTRACE = '/root/reference/vtrace.py'
REFERENCE_EXPERIMENT = '/root/reference/experiment.py'
_TF_NODES = (ast.With, ast.withitem, ast.Module, ast.Dict, ast.Invert)
_VTRACE_CALLS = frozenset({'log_probs_from_logits_and_actions', 'from_importance_weights', 'from_logits',
                           'VTraceFromLogitsReturns', 'VTraceReturns'})
_LOSS_FUNCTIONS = ('compute_baseline_loss', 'compute_entropy_loss', 'compute_policy_gradient_loss')
_GLUE_ARGS = ('FLAGS', 'agent_outputs', 'env_outputs', 'learner_outputs')
_GLUE_RESULTS = ('total_loss', 'vtrace_returns', 'clipped_rewards', 'discounts', 'learner_outputs',
                 'agent_outputs', 'bootstrap_value')


def tf_path_available():
  return os.path.exists(REFERENCE_VTRACE) and os.path.exists(REFERENCE_EXPERIMENT)


def _assigned_names(node):
  targets = node.targets if isinstance(node, ast.Assign) else [node.target]
  return {t.id for t in targets if isinstance(t, ast.Name)}


def lift_tf_path(tf):
  """Executes the reference's TF code for the path with `tf` (oracle.tf_shim).

  Returns a namespace with ``vtrace`` (from_logits, from_importance_weights,
  log_probs_from_logits_and_actions as vtrace.py defines them), the three
  ``compute_*_loss`` functions of experiment.py and ``learner_loss(FLAGS,
  agent_outputs, env_outputs, learner_outputs)`` = lines 365-407 of
  build_learner on the [T+1] structures (bootstrap value, the T+1 -> T shifts,
  reward clipping, discounts, V-trace, weighted total), returning the locals
  named in _GLUE_RESULTS.
  """
  builtins_allowed = {k: getattr(__import__('builtins'), k) for k in
                      ('range', 'reversed', 'len', 'float', 'int', 'zip', 'enumerate')}

  # vtrace.py: everything but its import statements.
  with open(REFERENCE_VTRACE, 'r') as f:
    tree = ast.parse(f.read(), filename=REFERENCE_VTRACE)
  body = [n for n in tree.body if not isinstance(n, (ast.Import, ast.ImportFrom))]
  module = ast.Module(body=body, type_ignores=[])
  _check_tree(module, _TF_NODES, _VTRACE_CALLS, frozenset({'_asdict'}))
  vt_ns = {'tf': tf, 'collections': collections, '__builtins__': builtins_allowed}
  exec(compile(module, REFERENCE_VTRACE, 'exec'), vt_ns)  # pylint: disable=exec-used
  vtrace_mod = types.SimpleNamespace(**{k: v for k, v in vt_ns.items() if not k.startswith('__')})

  with open(REFERENCE_EXPERIMENT, 'r') as f:
    tree = ast.parse(f.read(), filename=REFERENCE_EXPERIMENT)
  losses = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in _LOSS_FUNCTIONS]
  if len(losses) != len(_LOSS_FUNCTIONS):
    raise RuntimeError('experiment.py lacks one of %s' % (_LOSS_FUNCTIONS,))
  learner = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == 'build_learner']
  if len(learner) != 1:
    raise RuntimeError('experiment.py has no build_learner')
  glue = []
  for node in learner[0].body:
    if isinstance(node, ast.Assign) and _assigned_names(node) == {'bootstrap_value'}:
      glue.append(node)                                        # :365
    elif isinstance(node, ast.Assign) and 'map_structure' in ast.dump(node.value):
      glue.append(node)                                        # :371-375, the T+1 -> T shifts
    elif isinstance(node, ast.If) and 'reward_clipping' in ast.dump(node.test):
      glue.append(node)                                        # :377-382
    elif isinstance(node, ast.Assign) and _assigned_names(node) == {'discounts'}:
      glue.append(node)                                        # :384
    elif isinstance(node, ast.With) and 'from_logits' in ast.dump(node):
      glue.append(node)                                        # :389-397
    elif isinstance(node, (ast.Assign, ast.AugAssign)) and _assigned_names(node) == {'total_loss'}:
      glue.append(node)                                        # :401-407
  kinds = [type(n).__name__ for n in glue]
  if kinds != ['Assign'] * 4 + ['If', 'Assign', 'With', 'Assign', 'AugAssign', 'AugAssign']:
    raise RuntimeError('build_learner does not look as expected: %s' % kinds)
  ret = ast.Return(value=ast.Tuple(
      elts=[ast.Name(id=name, ctx=ast.Load()) for name in _GLUE_RESULTS], ctx=ast.Load()))
  fn = ast.FunctionDef(
      name='learner_loss',
      args=ast.arguments(posonlyargs=[], args=[ast.arg(arg=a) for a in _GLUE_ARGS], kwonlyargs=[],
                         kw_defaults=[], defaults=[]),
      body=glue + [ret], decorator_list=[], type_params=[])
  module = ast.fix_missing_locations(ast.Module(body=losses + [fn], type_ignores=[]))
  _check_tree(module, _TF_NODES + (ast.Lambda,), frozenset(_LOSS_FUNCTIONS))
  ex_ns = {'tf': tf, 'vtrace': vtrace_mod, 'nest': tf.contrib.framework.nest,  # experiment.py:40
           '__builtins__': builtins_allowed}
  exec(compile(module, REFERENCE_EXPERIMENT, 'exec'), ex_ns)  # pylint: disable=exec-used
  out = {k: ex_ns[k] for k in _LOSS_FUNCTIONS + ('learner_loss',)}
  out['vtrace'] = vtrace_mod
  out['glue_lines'] = (glue[0].lineno, glue[-1].end_lineno)
  return types.SimpleNamespace(**out)
