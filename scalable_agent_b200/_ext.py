"""Loader of the optional C++ host layer (csrc_torch/vtrace_torch.cpp).

The extension holds no kernel: it is the PyTorch plumbing of `_ops.FusedLoss`
/ `_ops.FusedLearnerLoss` in C++ (argument checks, allocation, the call into
libvtrace_b200.so, the autograd node), an order of magnitude cheaper per call
than the same thing in Python.  It is built in-tree by
`__graft_entry__.build()` (`python -m scalable_agent_b200._ext`) into
`scalable_agent_b200/lib/torch_ext/`; when it is absent the Python layer runs
the same kernels, only with more host time per call.  `VTRACE_B200_NO_EXT=1`
switches it off.
"""
import ctypes
import importlib.util
import os
import sys
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
NAME = 'vtrace_torch_ext'
BUILD_DIR = os.path.join(_HERE, 'lib', 'torch_ext')
SO_PATH = os.path.join(BUILD_DIR, NAME + '.so')
SOURCE = os.path.join(_HERE, 'csrc_torch', 'vtrace_torch.cpp')

_ext = None
_tried = False
_lock = threading.Lock()


def build(verbose=False):
  """Compiles the extension in-tree (host C++ only; needs no GPU)."""
  from torch.utils import cpp_extension
  os.makedirs(BUILD_DIR, exist_ok=True)
  os.environ.setdefault('TORCH_CUDA_ARCH_LIST', '10.0a')   # no .cu here; silences the arch probe
  cpp_extension.load(
      name=NAME, sources=[SOURCE], build_directory=BUILD_DIR, with_cuda=True,
      extra_cflags=['-O2', '-std=c++17'], is_python_module=True, verbose=verbose)
  return SO_PATH


def _addr(fn):
  return ctypes.cast(fn, ctypes.c_void_p).value


def get():
  """The bound extension module, or None when it is not built / switched off."""
  global _ext, _tried
  if _tried:
    return _ext
  with _lock:
    if _tried:
      return _ext
    ext = None
    if not os.environ.get('VTRACE_B200_NO_EXT') and os.path.exists(SO_PATH):
      try:
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        spec = importlib.util.spec_from_file_location(NAME, SO_PATH)
        module = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(module)
        from scalable_agent_b200 import _lib
        lib = _lib.load()
        module.bind(_addr(lib.vtr_from_logits_fused_f32), _addr(lib.vtr_fused_workspace_bytes),
                    _addr(lib.vtr_scale_inplace_f32), _addr(lib.vtr_last_error))
        module.bind_importance_weights(_addr(lib.vtr_from_importance_weights_f32))
        ext = module
      except Exception as e:  # pylint: disable=broad-except
        sys.stderr.write('scalable_agent_b200: C++ host layer not usable (%s); '
                         'falling back to the Python layer\n' % (str(e).splitlines()[0][:200],))
        ext = None
    _ext, _tried = ext, True
  return _ext


if __name__ == '__main__':
  print(build(verbose='-v' in sys.argv))
