"""B200-native V-trace + IMPALA loss: a drop-in for the learner-side
off-policy correction of Zhehui-Huang/scalable_agent.

    from scalable_agent_b200 import vtrace          # vtrace.py drop-in
    from scalable_agent_b200 import experiment      # loss helpers + fused loss

Importing the package is cheap and works without a GPU; calling any op needs
libvtrace_b200.so (built in-tree) and a CUDA device.
"""
from scalable_agent_b200 import _lib  # noqa: F401

__version__ = '0.1.0'


def __getattr__(name):
  # Lazy: `vtrace` / `experiment` pull in torch custom-op registration.
  if name in ('vtrace', 'experiment', 'distributed'):
    import importlib
    return importlib.import_module('scalable_agent_b200.' + name)
  raise AttributeError(name)
