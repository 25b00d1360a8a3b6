"""A ~60-line pytree with JAX semantics (None is an empty node)."""
_REGISTRY = {}


def register_pytree_node(cls, flatten, unflatten):
  _REGISTRY[cls] = (flatten, unflatten)


class _Leaf:
  def __repr__(self):
    return '*'


_LEAF = _Leaf()


def _flatten(x, leaves):
  if x is None:
    return ('none',)
  t = type(x)
  if t in _REGISTRY:
    children, aux = _REGISTRY[t][0](x)
    return ('reg', t, aux, tuple(_flatten(c, leaves) for c in children))
  if t is list or t is tuple:
    return ('seq', t, tuple(_flatten(c, leaves) for c in x))
  if t is dict:
    keys = sorted(x.keys())
    return ('dict', tuple(keys), tuple(_flatten(x[k], leaves) for k in keys))
  leaves.append(x)
  return _LEAF


def _unflatten(td, it):
  if td is _LEAF:
    return next(it)
  kind = td[0]
  if kind == 'none':
    return None
  if kind == 'reg':
    _, t, aux, ch = td
    return _REGISTRY[t][1](aux, [_unflatten(c, it) for c in ch])
  if kind == 'seq':
    _, t, ch = td
    return t(_unflatten(c, it) for c in ch)
  if kind == 'dict':
    _, keys, ch = td
    return {k: _unflatten(c, it) for k, c in zip(keys, ch)}
  raise TypeError(kind)


def tree_flatten(x):
  leaves = []
  td = _flatten(x, leaves)
  return leaves, td


def tree_unflatten(td, leaves):
  return _unflatten(td, iter(leaves))


def tree_leaves(x):
  return tree_flatten(x)[0]


def tree_map(f, tree, *rest):
  leaves, td = tree_flatten(tree)
  others = [tree_flatten(r)[0] for r in rest]
  for o in others:
    if len(o) != len(leaves):
      raise ValueError('tree_map: tree structure mismatch')
  return tree_unflatten(td, [f(*xs) for xs in zip(leaves, *others)])
