"""brax_b200 -- B200-native (sm_100a) implementation of Brax v1 `System.step`.

Only the hot path of the reference is here: the fused
substep loop as hand-written CUDA behind a C ABI (include/brax_b200.h), plus
the thin host-side mirror of the reference's `System.step(qp, act)` interface.
"""
__version__ = '0.1.0'


def __getattr__(name):
  # `brax_b200.System`, `.QP`, `.AutoReset`, `.load_system` ...: resolved on first use so that
  # importing the package (e.g. for `brax_b200.build`) does not need the built library
  if name in ('System', 'QP', 'Info', 'JointState', 'AutoReset', 'HostBuffers', 'load_system'):
    from brax_b200 import system
    return getattr(system, name)
  raise AttributeError('module %r has no attribute %r' % (__name__, name))
