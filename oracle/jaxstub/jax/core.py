"""jax.core stand-in: jumpy.py:39 reads core.cur_sublevel().level."""


class _Sublevel:
  level = 0


def cur_sublevel():
  return _Sublevel()
