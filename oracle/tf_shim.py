"""A stand-in for the TensorFlow 1.x names the reference's hot path uses.  TEST INFRASTRUCTURE ONLY.

TensorFlow 1.x cannot be installed in this image, so the reference's own
``vtrace.py`` and the loss code of ``experiment.py`` cannot run as they are.
They use a small, elementary part of the TF API, though.  This module provides
exactly those names on top of torch CPU tensors (float64 by default, float32
on request), with autograd, so that the REFERENCE'S OWN SOURCE TEXT can be
executed here (``oracle/lift_reference.py``) and its outputs and gradients
frozen as fixtures (``tests/golden/make_golden_losses.py``).

What that pins and what it does not: the composition - which quantities are
formed, clipped, shifted, stopped, summed and weighted, in which order - is
the reference's, line for line.  The primitives below are restated from the
TF 1.x API documentation, one line each:

=============================================  =================================
reference use (file:line)                      here
=============================================  =================================
tf.convert_to_tensor(x, dtype)  vtrace.py:62   torch.as_tensor, working dtype
x.shape.assert_has_rank / .ndims  :64,:231     Shape (a tuple subclass)
tf.name_scope / tf.device      :85, exp.py:389 no-op context managers
tf.exp / tf.minimum / tf.add   :245-264        torch.exp / minimum / add
tf.concat / tf.expand_dims     :252,:266       torch.cat / unsqueeze
tf.zeros_like                  :259            torch.zeros_like
tf.scan(reverse=True, back_prop=False) :260    a Python loop from the back
tf.stop_gradient               :279            Tensor.detach
tf.nn.sparse_softmax_cross_entropy_with_logits -log_softmax gathered at labels
tf.nn.softmax / log_softmax    exp.py:332-333  torch.softmax / log_softmax
tf.reduce_sum / tf.square      exp.py:328      Tensor.sum / torch.square
tf.clip_by_value / tanh / where exp.py:378-382 torch.clamp / tanh / where
tf.to_float                    exp.py:384      cast to the working dtype
=============================================  =================================

Nothing here is imported by the product.
"""
import contextlib
import types

import numpy as np
import torch

_WORKING = [torch.float64]


def set_precision(name):
  """'float64' (the pin) or 'float32' (how the reference itself computes)."""
  _WORKING[0] = {'float64': torch.float64, 'float32': torch.float32}[name]


class Shape(tuple):
  """TensorShape as far as the reference uses it."""

  @property
  def ndims(self):
    return len(self)

  def assert_has_rank(self, rank):
    if len(self) != rank:
      raise ValueError('Shape %s must have rank %d' % (tuple(self), rank))


class Tensor(torch.Tensor):
  """torch.Tensor whose .shape answers like a TensorShape."""

  @property
  def shape(self):  # pylint: disable=invalid-overridden-method
    return Shape(self.size())


def _wrap(x):
  return x.as_subclass(Tensor) if isinstance(x, torch.Tensor) else x


class _DType(object):
  def __init__(self, floating):
    self.floating = floating


float32 = _DType(True)
int32 = _DType(False)


def convert_to_tensor(value, dtype=None, name=None):
  del name
  want = None
  if dtype is not None:
    want = _WORKING[0] if dtype.floating else torch.int64
  if isinstance(value, torch.Tensor):
    return _wrap(value if want is None or value.dtype == want else value.to(want))
  out = torch.as_tensor(np.asarray(value))  # through numpy: a Python float stays a double
  if want is None and out.is_floating_point():
    want = _WORKING[0]
  return _wrap(out if want is None else out.to(want))


def _t(x):
  return x if isinstance(x, torch.Tensor) else convert_to_tensor(x, float32)


@contextlib.contextmanager
def name_scope(name=None, default_name=None, values=None):
  del name, default_name, values
  yield


@contextlib.contextmanager
def device(name):
  del name
  yield


def exp(x, name=None):
  return _wrap(torch.exp(_t(x)))


def tanh(x, name=None):
  return _wrap(torch.tanh(_t(x)))


def square(x, name=None):
  return _wrap(torch.square(_t(x)))


def minimum(x, y, name=None):
  return _wrap(torch.minimum(_t(x), _t(y)))


def add(x, y, name=None):
  return _wrap(torch.add(_t(x), _t(y)))


def concat(values, axis, name=None):
  return _wrap(torch.cat([_t(v) for v in values], dim=axis))


def expand_dims(x, axis, name=None):
  return _wrap(torch.unsqueeze(_t(x), axis))


def zeros_like(x, name=None):
  return _wrap(torch.zeros_like(_t(x)))


def stop_gradient(x, name=None):
  return _wrap(_t(x).detach())


def reduce_sum(x, axis=None, name=None):
  x = _t(x)
  return _wrap(x.sum() if axis is None else x.sum(dim=axis))


def clip_by_value(x, lo, hi, name=None):
  return _wrap(torch.clamp(_t(x), min=lo, max=hi))


def where(condition, x, y, name=None):
  return _wrap(torch.where(condition, _t(x), _t(y)))


def to_float(x, name=None):
  return _wrap(_t(x).to(_WORKING[0]))


def scan(fn, elems, initializer, parallel_iterations=1, back_prop=True, reverse=False, name=None):
  """tf.scan over the leading axis of every tensor of the tuple `elems`;
  returns the stacked accumulators, in the order of the inputs."""
  del parallel_iterations, name
  single = not isinstance(elems, (tuple, list))
  seqs = [_t(e) for e in ([elems] if single else elems)]
  n = seqs[0].size(0)
  order = range(n - 1, -1, -1) if reverse else range(n)
  acc = _t(initializer)
  out = [None] * n
  for i in order:
    item = seqs[0][i] if single else tuple(s[i] for s in seqs)
    acc = fn(acc, item)
    out[i] = acc
  stacked = torch.stack(out, dim=0)
  return _wrap(stacked if back_prop else stacked.detach())


def _sparse_softmax_cross_entropy_with_logits(_sentinel=None, labels=None, logits=None, name=None):
  del _sentinel, name
  logits = _t(logits)
  labels = torch.as_tensor(labels).to(torch.int64)
  lsm = torch.log_softmax(logits, dim=-1)
  return _wrap(-torch.gather(lsm, -1, labels.unsqueeze(-1)).squeeze(-1))


nn = types.SimpleNamespace(
    sparse_softmax_cross_entropy_with_logits=_sparse_softmax_cross_entropy_with_logits,
    softmax=lambda logits, axis=-1, name=None: _wrap(torch.softmax(_t(logits), dim=axis)),
    log_softmax=lambda logits, axis=-1, name=None: _wrap(torch.log_softmax(_t(logits), dim=axis)))


def _map_structure(fn, *structures):
  first = structures[0]
  if isinstance(first, tuple) and hasattr(first, '_fields'):
    return type(first)(*[_map_structure(fn, *parts) for parts in zip(*structures)])
  if isinstance(first, (tuple, list)):
    return type(first)(_map_structure(fn, *parts) for parts in zip(*structures))
  return fn(*structures)


# vtrace.py:34 binds `nest = tf.contrib.framework.nest` at import time.
contrib = types.SimpleNamespace(
    framework=types.SimpleNamespace(nest=types.SimpleNamespace(map_structure=_map_structure)))
