"""ctypes binding of libvtrace_b200.so (the C ABI in include/vtrace_b200.h).

There is no CPU fallback and no alternative backend: if the shared library is
missing or does not export a declared symbol, importing an op raises.
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# VTRACE_B200_LIB points at another build of the same ABI (A/B experiments).
LIB_PATH = os.environ.get('VTRACE_B200_LIB') or os.path.join(
    _HERE, 'lib', 'libvtrace_b200.so')

VTR_OK = 0
VTR_MAX_ACTIONS = 256
FLAG_NO_CLIP_RHO = 1
FLAG_NO_CLIP_PG_RHO = 2
REWARD_CLIP_MODES = {None: 0, 'none': 0, 'abs_one': 1, 'soft_asymmetric': 2}
PATHS = {None: 0, 'auto': 0, 'column': 1, 'sm_tile': 2, 'item': 3, 'chunk': 4}


class FusedParams(ctypes.Structure):
  """Mirror of `struct vtr_fused_params` (include/vtrace_b200.h)."""
  _fields_ = [
      ('T', ctypes.c_int32), ('B', ctypes.c_int32), ('A', ctypes.c_int32),
      ('flags', ctypes.c_uint32),
      ('behaviour_policy_logits', ctypes.c_void_p),
      ('target_policy_logits', ctypes.c_void_p),
      ('actions', ctypes.c_void_p),
      ('discounts', ctypes.c_void_p),
      ('rewards', ctypes.c_void_p),
      ('values', ctypes.c_void_p),
      ('bootstrap_value', ctypes.c_void_p),
      ('done', ctypes.c_void_p),
      ('clip_rho_threshold', ctypes.c_float),
      ('clip_pg_rho_threshold', ctypes.c_float),
      ('baseline_cost', ctypes.c_float),
      ('entropy_cost', ctypes.c_float),
      ('discounting', ctypes.c_float),
      ('reward_clip_mode', ctypes.c_int32),
      ('path', ctypes.c_int32),
      ('reserved', ctypes.c_int32),
      ('vs', ctypes.c_void_p),
      ('pg_advantages', ctypes.c_void_p),
      ('log_rhos', ctypes.c_void_p),
      ('behaviour_action_log_probs', ctypes.c_void_p),
      ('target_action_log_probs', ctypes.c_void_p),
      ('d_target_logits', ctypes.c_void_p),
      ('d_values', ctypes.c_void_p),
      ('losses', ctypes.c_void_p),
      ('losses_f64', ctypes.c_void_p),
      ('exchange_peer_buffers', ctypes.POINTER(ctypes.c_void_p)),
      ('exchange_rank', ctypes.c_int32),
      ('exchange_world', ctypes.c_int32),
      ('exchange_seq', ctypes.c_uint32),
      ('exchange_lag', ctypes.c_uint32),
      ('exchange_sums', ctypes.c_void_p),
      ('total_loss', ctypes.c_void_p),
  ]


_P = ctypes.c_void_p
_I32 = ctypes.c_int32
_I64 = ctypes.c_int64
_U32 = ctypes.c_uint32
_F = ctypes.c_float
_SZ = ctypes.c_size_t

# name -> (restype, argtypes); every symbol include/vtrace_b200.h declares.
SIGNATURES = {
    'vtr_version': (ctypes.c_int, []),
    'vtr_last_error': (ctypes.c_char_p, []),
    'vtr_fused_workspace_bytes': (_SZ, [_I32, _I32, _I32]),
    'vtr_from_logits_fused_f32':
        (ctypes.c_int, [ctypes.POINTER(FusedParams), _P, _SZ, _P]),
    'vtr_from_logits_fused_noalloc_f32':
        (ctypes.c_int, [ctypes.POINTER(FusedParams), _P]),
    'vtr_fused_describe':
        (ctypes.c_int, [ctypes.POINTER(FusedParams), ctypes.c_int,
                        ctypes.POINTER(ctypes.c_int), ctypes.c_char_p, _SZ]),
    'vtr_reduce_losses': (ctypes.c_int, [_P, _I32, _F, _F, _P, _P, _P]),
    'vtr_from_importance_weights_f32':
        (ctypes.c_int, [_P, _P, _P, _P, _P, _I32, _I32, _I32, _U32, _F, _F,
                        _U32, _P, _P, _P]),
    'vtr_log_probs_from_logits_and_actions_f32':
        (ctypes.c_int, [_P, _P, _I64, _I32, _P, _P]),
    'vtr_loss_workspace_bytes': (_SZ, [_I64]),
    'vtr_policy_gradient_loss_f32':
        (ctypes.c_int, [_P, _P, _P, _I64, _I32, _P, _P, _SZ, _P]),
    'vtr_policy_gradient_loss_bwd_f32':
        (ctypes.c_int, [_P, _P, _P, _P, _I64, _I32, _P, _P]),
    'vtr_entropy_loss_f32': (ctypes.c_int, [_P, _I64, _I32, _P, _P, _SZ, _P]),
    'vtr_entropy_loss_bwd_f32': (ctypes.c_int, [_P, _P, _I64, _I32, _P, _P]),
    'vtr_baseline_loss_f32': (ctypes.c_int, [_P, _I64, _P, _P, _SZ, _P]),
    'vtr_baseline_loss_bwd_f32': (ctypes.c_int, [_P, _P, _I64, _P, _P]),
    'vtr_scale_inplace_f32': (ctypes.c_int, [_P, _P, _I64, _P]),
    'vtr_exchange_buffer_bytes': (_SZ, [_I32]),
    'vtr_exchange_publish_f32':
        (ctypes.c_int, [_P, _I32, ctypes.POINTER(ctypes.c_void_p), _I32, _I32, _U32, _P]),
    'vtr_exchange_collect_f32': (ctypes.c_int, [_P, _I32, _I32, _U32, _P, _P]),
    'vtr_polynomial_decay':
        (ctypes.c_double, [ctypes.c_double, _I64, _I64, ctypes.c_double, ctypes.c_double]),
    'vtr_rmsprop_f32': (ctypes.c_int, [_P, _P, _P, _P, _I64, _F, _F, _F, _F, _P]),
    'vtr_rmsprop_slice':
        (ctypes.c_int, [_I64, _I32, _I32, ctypes.POINTER(_I64), ctypes.POINTER(_I64)]),
    'vtr_rmsprop_sync_bytes': (_SZ, []),
    'vtr_rmsprop_allreduce_f32':
        (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_void_p),
                        ctypes.POINTER(ctypes.c_void_p), _P, _P, _I64, _I32, _I32, _U32,
                        _F, _F, _F, _F, _P, _P]),
    'vtr_policy_head_packed_bytes': (_SZ, [_I32, _I32]),
    'vtr_policy_head_prepare_f32': (ctypes.c_int, [_P, _P, _P, _P, _I32, _I32, _P, _P]),
    'vtr_policy_head_f32': (ctypes.c_int, [_P, _P, _I64, _I32, _I32, _P, _P, _P]),
    'vtr_set_peer_timeout_ms': (None, [ctypes.c_uint64]),
    'vtr_get_peer_timeout_ms': (ctypes.c_uint64, []),
}

_lib = None
_lock = threading.Lock()


class LibraryMissingError(ImportError):
  pass


def load():
  """Loads the library once; raises loudly if it is not built."""
  global _lib
  if _lib is not None:
    return _lib
  with _lock:
    if _lib is not None:
      return _lib
    if not os.path.exists(LIB_PATH):
      raise LibraryMissingError(
          '%s not found. Build it with `make -C scalable_agent_b200/csrc -j` '
          '(or `python -c "import __graft_entry__ as g; g.build()"`). '
          'scalable_agent_b200 has no CPU or PyTorch fallback.' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SIGNATURES.items():
      try:
        fn = getattr(lib, name)
      except AttributeError:
        raise LibraryMissingError(
            '%s does not export %s; rebuild it.' % (LIB_PATH, name))
      fn.restype = restype
      fn.argtypes = argtypes
    _lib = lib
  return _lib


def last_error():
  return load().vtr_last_error().decode('utf-8', 'replace')


def check(rc, what):
  """Maps the C return convention to Python exceptions."""
  if rc == VTR_OK:
    return
  message = '%s failed (%d): %s' % (what, rc, last_error())
  if rc < 0:
    raise ValueError(message)
  raise RuntimeError(message)
