"""Drop-in of the CUDA path into a constructed reference `brax.System`: overrides `step` only.

    import brax, brax_b200.shim
    sys = brax.System(config)            # or env.sys of any reference env
    brax_b200.shim.install(sys)          # sys.step(qp, act) -> (brax.QP, brax.Info) through bx_step[_host]

Everything else of the reference -- `default_qp`, `info`, the envs' `_get_obs`, wrappers, agents -- keeps running
as it is and sees the reference's own `QP` / `Info` types (brax/physics/base.py:75-153).  The arrays go through
`brax_b200.System.step`: numpy arrays take the host path (`bx_step_host`), CUDA arrays of any `__dlpack__` /
`__cuda_array_interface__` producer (a jax.Array on the GPU) are read in place and the results come back as torch
CUDA tensors, which `jax.dlpack.from_dlpack` adopts without a copy.  Under `jax.jit` / `vmap` / `lax.scan` the
call has to be an XLA custom call instead: that is `brax_b200.jax_ffi` (source only on this box: no jaxlib).

tests/test_extract_reference.py installs this shim into the reference's own `Ant` env (under the numpy stand-in
for jax) with the float64 oracle as the backend and checks that `Env.step` returns what the unpatched reference
returns -- the adapter's plumbing; the CUDA backend itself is what tests/test_gpu_*.py cover.
"""
import numpy as np

from brax_b200 import system as _system


def make_step(brax_system, backend=None, name: str = 'system', act_size=None, device: int = 0):
  """`step(qp, act) -> (qp, info)` with the reference's types for a constructed reference `brax.System`.
  backend: an object with `System.step`'s signature (default: `brax_b200.System.from_brax(brax_system)`)."""
  from brax.physics.base import Info, P, QP
  b200 = backend if backend is not None else _system.System.from_brax(brax_system, name, act_size=act_size, device=device)

  def host(x):
    return x if (_system._is_torch(x) or hasattr(x, '__dlpack__') and not isinstance(x, np.ndarray)) else np.asarray(x, np.float32)

  def step(qp, act):
    q = _system.QP(host(qp.pos), host(qp.rot), host(qp.vel), host(qp.ang))
    a = host(act)
    if isinstance(a, np.ndarray) and a.shape[-1] < b200.act_size:
      # systems whose env feeds a shorter act vector than the descriptor's (grasp's palm-target slots)
      a = np.concatenate([a, np.zeros(a.shape[:-1] + (b200.act_size - a.shape[-1],), np.float32)], -1)
    out, info = b200.step(q, a)
    return (QP(pos=out.pos, rot=out.rot, vel=out.vel, ang=out.ang),
            Info(contact=P(vel=info.contact.vel, ang=info.contact.ang), joint=P(vel=info.joint.vel, ang=info.joint.ang),
                 actuator=P(vel=info.actuator.vel, ang=info.actuator.ang), contact_pos=info.contact_pos,
                 contact_normal=info.contact_normal, contact_penetration=info.contact_penetration))

  step._b200 = b200       # keeps the device tables alive
  return step


def install(brax_system, **kw):
  """Overrides `step` on this `brax.System` instance only (envs, wrappers and agents unchanged)."""
  object.__setattr__(brax_system, 'step', make_step(brax_system, **kw))
  return brax_system
