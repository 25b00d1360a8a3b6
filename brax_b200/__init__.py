"""brax_b200 -- B200-native (sm_100a) implementation of Brax v1 `System.step`.

Only the hot path of the reference is here: the fused
substep loop as hand-written CUDA behind a C ABI (include/brax_b200.h), plus
the thin host-side mirror of the reference's `System.step(qp, act)` interface.
"""
__version__ = '0.1.0'
