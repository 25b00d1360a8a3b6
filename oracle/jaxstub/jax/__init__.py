"""TEST INFRASTRUCTURE ONLY -- a minimal stand-in for the `jax` package.

It exists so that the *unmodified* reference (`/root/reference/brax`) can be
imported in a container that has no JAX, and run on its own numpy backend
(`brax/jumpy.py:42-48` dispatches every op to numpy when no argument is a JAX
array; `jumpy.py:54-91,99-119` unroll vmap/scan into Python loops outside jit).
Nothing in the product imports this.  Only `oracle/gen_golden.py` and
`tools/gen_systems.py` put this directory on sys.path, and only in the build
container (the GPU box has no /root/reference).

Semantics that matter:
  * tree_util treats None as an EMPTY NODE, not a leaf (colliders.py:133,220
    rely on dp_b/dq_b staying None through vmap/tree_map).
  * registered classes use (children, aux) / unflatten(aux, children).
  * core.cur_sublevel().level == 0 and config.jax_disable_jit == False, so
    jumpy._in_jit() is False and the numpy path is taken.
"""
from . import tree_util, core, numpy, interpreters  # noqa: F401
from .interpreters import batching  # noqa: F401


class _Config:
  jax_disable_jit = False

  def update(self, *a, **k):
    pass


config = _Config()


class custom_jvp:  # jumpy.py:308-340
  def __init__(self, fn):
    self.fn = fn

  def __call__(self, *a, **k):
    return self.fn(*a, **k)

  def defjvp(self, fn):
    return fn


def jit(fn, *a, **k):
  return fn


def vmap(*a, **k):
  raise NotImplementedError('jax stub: vmap is only reachable inside jit')


class _Lax:
  def __getattr__(self, name):
    raise NotImplementedError('jax stub: lax.%s' % name)


lax = _Lax()


class _Ops:
  def __getattr__(self, name):
    raise NotImplementedError('jax stub: ops.%s' % name)


ops = _Ops()
