"""Sharding of a batch of environments over ranks (one process per GPU).

Environments are independent (`System.step` has no cross-env term: every op
of the reference is under `vmap`, envs/wrappers.py:70,80), so a batch shards
into contiguous blocks with NO per-step collective -- the same sharding the
reference's trainers use under `jax.pmap` (agents/ppo/train.py:280-282).  The
only communication is one reduction of a small stats vector after the timed
loop.  Backend: nccl on GPUs, gloo in the CPU tests.
"""
from typing import Sequence, Tuple


def shard_range(total_envs: int, world: int, rank: int) -> Tuple[int, int]:
  """[lo, hi) of the contiguous env block of `rank`; sizes differ by <= 1."""
  if world < 1 or not 0 <= rank < world:
    raise ValueError('bad world/rank %d/%d' % (world, rank))
  if total_envs < 0:
    raise ValueError('total_envs must be >= 0')
  base, rem = divmod(total_envs, world)
  lo = rank * base + min(rank, rem)
  return lo, lo + base + (1 if rank < rem else 0)


def shard_sizes(total_envs: int, world: int) -> Sequence[int]:
  return [shard_range(total_envs, world, r)[1] - shard_range(total_envs, world, r)[0]
          for r in range(world)]


def reduce_stats(values: Sequence[float], ops: Sequence[str], device=None):
  """All-reduces a small vector of per-rank statistics.

  ops[i] in {'max', 'sum', 'min'}: timings take 'max' (a multi-GPU number is
  the slowest rank's), counts take 'sum'.  Without an initialised process
  group (single process) the values are returned unchanged."""
  import torch
  import torch.distributed as dist
  if len(values) != len(ops):
    raise ValueError('values/ops length mismatch')
  if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
    return [float(v) for v in values]
  out = [0.0] * len(values)
  red = {'max': dist.ReduceOp.MAX, 'sum': dist.ReduceOp.SUM, 'min': dist.ReduceOp.MIN}
  for kind, op in red.items():
    idx = [i for i, o in enumerate(ops) if o == kind]
    if not idx:
      continue
    t = torch.tensor([float(values[i]) for i in idx], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=op)
    for i, v in zip(idx, t.tolist()):
      out[i] = float(v)
  bad = [o for o in ops if o not in red]
  if bad:
    raise ValueError('unknown reduce op %r' % bad[0])
  return out
