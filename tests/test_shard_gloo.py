"""N>1 path on CPU: env sharding + the final stats reduction over gloo (world 2).

The data path has no collective (envs are independent), so what needs covering
is (1) the partition: every env owned by exactly one rank, blocks contiguous;
(2) the one reduction bench.py does after the timed loop (max of timings, sum
of counts); (3) that stepping a shard gives the same values as stepping the
whole batch -- checked here with the oracle standing in for the kernel (the
CHECKER role: there is no GPU in this container).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from brax_b200 import shard
from tests import util


def test_shard_range_partitions_everything():
  for total in (0, 1, 7, 128, 8192, 8193, 1_000_003):
    for world in (1, 2, 3, 4, 8):
      edges = [shard.shard_range(total, world, r) for r in range(world)]
      assert edges[0][0] == 0 and edges[-1][1] == total
      for (a, b), (c, d) in zip(edges, edges[1:]):
        assert b == c and a <= b
      sizes = shard.shard_sizes(total, world)
      assert sum(sizes) == total and max(sizes) - min(sizes) <= 1
  with pytest.raises(ValueError):
    shard.shard_range(10, 2, 2)


def _free_port():
  s = socket.socket()
  s.bind(('127.0.0.1', 0))
  p = s.getsockname()[1]
  s.close()
  return p


def _worker(rank, world, port, total, out_dir):
  os.environ['MASTER_ADDR'] = '127.0.0.1'
  os.environ['MASTER_PORT'] = str(port)
  dist.init_process_group('gloo', rank=rank, world_size=world)
  try:
    from oracle.oracle import Oracle
    g = util.load_golden('ant')
    lo, hi = shard.shard_range(total, world, rank)
    o = Oracle(util.load_desc('ant'), np.float32)
    out = o.step(g['pos'][lo:hi], g['rot'][lo:hi], g['vel'][lo:hi], g['ang'][lo:hi],
                 g['act'][lo:hi], num_threads=1)
    np.save(os.path.join(out_dir, 'pos_%d.npy' % rank), out['pos'])
    # timings -> max, env counts -> sum (what bench.py reduces)
    red = shard.reduce_stats([1.0 + rank, hi - lo, 5.0 - rank], ['max', 'sum', 'min'])
    np.save(os.path.join(out_dir, 'red_%d.npy' % rank), np.asarray(red))
  finally:
    dist.destroy_process_group()


def test_two_ranks_gloo(tmp_path):
  world, total = 2, 21
  port = _free_port()
  mp.spawn(_worker, args=(world, port, total, str(tmp_path)), nprocs=world, join=True)
  from oracle.oracle import Oracle
  g = util.load_golden('ant')
  full = Oracle(util.load_desc('ant'), np.float32).step(
      g['pos'][:total], g['rot'][:total], g['vel'][:total], g['ang'][:total], g['act'][:total])
  parts = [np.load(tmp_path / ('pos_%d.npy' % r)) for r in range(world)]
  np.testing.assert_array_equal(np.concatenate(parts), full['pos'])
  for r in range(world):
    red = np.load(tmp_path / ('red_%d.npy' % r))
    np.testing.assert_array_equal(red, [2.0, float(total), 4.0])


def test_reduce_stats_single_process():
  assert shard.reduce_stats([3.0, 4.0], ['max', 'sum']) == [3.0, 4.0]
  with pytest.raises(ValueError):
    shard.reduce_stats([1.0], ['max', 'sum'])
