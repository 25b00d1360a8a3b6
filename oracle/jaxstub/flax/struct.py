"""flax.struct stand-in (base.py:25-28, envs/env.py:28-37)."""
import dataclasses
import jax

field = dataclasses.field


def dataclass(cls):
  cls = dataclasses.dataclass(cls)
  names = [f.name for f in dataclasses.fields(cls)]

  def flatten(obj):
    return [getattr(obj, n) for n in names], None

  def unflatten(_, children):
    obj = object.__new__(cls)
    for n, c in zip(names, children):
      object.__setattr__(obj, n, c)
    return obj

  jax.tree_util.register_pytree_node(cls, flatten, unflatten)
  cls.replace = lambda self, **kw: dataclasses.replace(self, **kw)
  return cls
