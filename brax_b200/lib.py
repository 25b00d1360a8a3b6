"""ctypes binding of libbrax_b200.so (include/brax_b200.h).

The product path fails loudly when the CUDA extension is missing: there is no
CPU fallback anywhere in brax_b200.
"""
import ctypes as C
import os

from brax_b200 import desc as _desc

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('BRAX_B200_LIB') or os.path.join(HERE, 'libbrax_b200.so')

SYMBOLS = ('bx_system_create', 'bx_system_destroy', 'bx_system_num_contacts',
           'bx_system_num_bodies', 'bx_system_act_size', 'bx_system_obs_size',
           'bx_system_num_metrics', 'bx_step',
           'bx_step_host', 'bx_rollout', 'bx_system_configure',
           'bx_system_autotune', 'bx_system_threads', 'bx_host_transfer_bytes',
           'bx_host_alloc', 'bx_host_free', 'bx_desc_tile_bytes',
           'bx_launch_count', 'bx_last_error', 'bx_system_num_joint_dof', 'bx_default_qp', 'bx_default_qp_offsets',
           'bx_step_host_static', 'bx_static_scratch_bytes', 'bx_step_host_async', 'bx_host_wait', 'bx_state_create', 'bx_state_destroy',
           'bx_state_num_envs', 'bx_state_set', 'bx_state_get', 'bx_state_rollout')

_lib = None


def load():
  """Loads the shared library (once).  Raises if it has not been built."""
  global _lib
  if _lib is not None:
    return _lib
  if not os.path.exists(LIB_PATH):
    raise RuntimeError(
        'brax_b200: %s is missing -- run `python -m brax_b200.build` '
        '(or __graft_entry__.build()); there is no CPU fallback.' % LIB_PATH)
  lib = C.CDLL(LIB_PATH)
  P = C.c_void_p
  lib.bx_system_create.argtypes = [C.POINTER(_desc.CSystemDesc), C.c_int,
                                   C.POINTER(P)]
  lib.bx_system_create.restype = C.c_int
  lib.bx_system_destroy.argtypes = [P]
  lib.bx_system_destroy.restype = None
  for f in ('bx_system_num_contacts', 'bx_system_num_bodies',
            'bx_system_act_size', 'bx_system_obs_size', 'bx_system_num_metrics',
            'bx_system_num_joint_dof'):
    getattr(lib, f).argtypes = [P]
    getattr(lib, f).restype = C.c_int
  lib.bx_default_qp.argtypes = [P, C.c_int64, P, P, P, P, P, P, P]
  lib.bx_default_qp.restype = C.c_int
  lib.bx_default_qp_offsets.argtypes = [P, C.c_int64, P, P, P, P, P, P, P, P]
  lib.bx_default_qp_offsets.restype = C.c_int
  lib.bx_step.argtypes = [P, C.c_int64, C.POINTER(_desc.CStepIO), P]
  lib.bx_step.restype = C.c_int
  lib.bx_step_host.argtypes = [P, C.c_int64, C.POINTER(_desc.CStepIO)]
  lib.bx_step_host.restype = C.c_int
  lib.bx_step_host_static.argtypes = [P, C.c_int64, C.POINTER(_desc.CStepIO), P]
  lib.bx_step_host_static.restype = C.c_int
  lib.bx_static_scratch_bytes.argtypes = []
  lib.bx_static_scratch_bytes.restype = C.c_int64
  lib.bx_step_host_async.argtypes = [P, C.c_int64, C.POINTER(_desc.CStepIO), C.c_int32]
  lib.bx_step_host_async.restype = C.c_int
  lib.bx_host_wait.argtypes = [P, C.c_int32]
  lib.bx_host_wait.restype = C.c_int
  lib.bx_state_create.argtypes = [P, C.c_int64, C.POINTER(P)]
  lib.bx_state_create.restype = C.c_int
  lib.bx_state_destroy.argtypes = [P]
  lib.bx_state_destroy.restype = None
  lib.bx_state_num_envs.argtypes = [P]
  lib.bx_state_num_envs.restype = C.c_int64
  lib.bx_state_set.argtypes = [P, P, P, P, P, P]
  lib.bx_state_set.restype = C.c_int
  lib.bx_state_get.argtypes = [P, P, P, P, P, P]
  lib.bx_state_get.restype = C.c_int
  lib.bx_state_rollout.argtypes = [P, C.c_int32, C.POINTER(_desc.CStepIO), P]
  lib.bx_state_rollout.restype = C.c_int
  lib.bx_rollout.argtypes = [P, C.c_int64, C.c_int32,
                             C.POINTER(_desc.CStepIO), P]
  lib.bx_rollout.restype = C.c_int
  lib.bx_system_configure.argtypes = [P, C.c_int32, C.c_int32]
  lib.bx_system_configure.restype = C.c_int
  lib.bx_system_autotune.argtypes = [P, C.c_int64, C.POINTER(_desc.CStepIO), P]
  lib.bx_system_autotune.restype = C.c_int
  lib.bx_system_threads.argtypes = [P]
  lib.bx_system_threads.restype = C.c_int
  lib.bx_host_transfer_bytes.argtypes = [P, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
  lib.bx_host_transfer_bytes.restype = C.c_int
  lib.bx_desc_tile_bytes.argtypes = [C.POINTER(_desc.CSystemDesc), C.POINTER(C.c_int64),
                                     C.POINTER(C.c_int32)]
  lib.bx_desc_tile_bytes.restype = C.c_int
  lib.bx_host_alloc.argtypes = [C.c_size_t]
  lib.bx_host_alloc.restype = C.c_void_p
  lib.bx_host_free.argtypes = [C.c_void_p]
  lib.bx_host_free.restype = None
  lib.bx_launch_count.argtypes = []
  lib.bx_launch_count.restype = C.c_int64
  lib.bx_last_error.argtypes = []
  lib.bx_last_error.restype = C.c_char_p
  _lib = lib
  return lib


def last_error() -> str:
  return load().bx_last_error().decode()


def launch_count() -> int:
  return int(load().bx_launch_count())
