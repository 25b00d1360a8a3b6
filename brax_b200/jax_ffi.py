"""JAX side of the boundary: `System.step` as an XLA FFI custom call (brax_b200/csrc/bx_ffi.cc).

    import brax, brax_b200.jax_ffi as bxj
    sys = brax.System(config)
    bxj.install(sys)            # sys.step(qp, act) -> ONE `bx_step` custom call, under jit / vmap / scan

**Unavailable on this box**: jax / jaxlib are not installed here or on the GPU box and there is no
network, so nothing in this module has been run; `brax_b200.build.build_ffi()` compiles the handler
only where `jax.ffi.include_dir()` resolves and says why it did not otherwise.  The module imports
without jax (everything jax-related happens inside the functions); it is the file a maintainer of the
reference adds, see INTEGRATION.md section 3.  The torch route (`brax_b200.System.step` on
`tensor.data_ptr()`, or any `__dlpack__` producer) is the built and tested one.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
FFI_LIB = os.path.join(HERE, 'libbrax_b200_ffi.so')
_registered = False


def register():
  """Registers the `bx_step` FFI target with XLA (once)."""
  global _registered
  if _registered:
    return
  import jax
  if not os.path.exists(FFI_LIB):
    raise RuntimeError('%s is missing: run brax_b200.build.build_ffi() where jaxlib is installed' % FFI_LIB)
  lib = ctypes.CDLL(FFI_LIB)
  jax.ffi.register_ffi_target('bx_step', jax.ffi.pycapsule(lib.bx_step_ffi), platform='CUDA')
  _registered = True


def make_step(brax_system, name: str = 'system', act_size=None, device: int = 0):
  """Returns `step(qp, act) -> (qp, info)` with the reference's types (brax/physics/base.py:75-153)
  for a constructed reference `brax.System`: one custom call; every leading dimension jp.vmap adds
  (envs/wrappers.py:70,80) folds into the batch of ONE launch (`vmap_method='expand_dims'`)."""
  import jax
  import jax.numpy as jnp
  from brax.physics.base import Info, P, QP
  from brax_b200.system import System
  register()
  b200 = System.from_brax(brax_system, name, act_size=act_size, device=device)
  n, k = b200.num_bodies, b200.num_contacts

  def step(qp, act):
    lead = qp.pos.shape[:-2]                      # () unbatched, (B,) under vmap
    f = lambda *s: jax.ShapeDtypeStruct(lead + s, jnp.float32)  # noqa: E731
    outs = jax.ffi.ffi_call(
        'bx_step', [f(n, 3), f(n, 4), f(n, 3), f(n, 3)] + [f(n, 3)] * 6 + [f(k, 3), f(k, 3), f(k)],
        vmap_method='expand_dims')(qp.pos, qp.rot, qp.vel, qp.ang, act, sys=jnp.int64(b200.handle))
    return (QP(*outs[:4]),
            Info(P(outs[4], outs[5]), P(outs[6], outs[7]), P(outs[8], outs[9]), outs[10], outs[11], outs[12]))

  step._b200 = b200       # keeps the device tables alive
  return step


def install(brax_system, **kw):
  """Overrides `step` on this `brax.System` instance only (envs, wrappers and agents unchanged)."""
  object.__setattr__(brax_system, 'step', make_step(brax_system, **kw))
  return brax_system
