"""Parity report (GPU box): the numbers behind the tolerance policy of tests/test_gpu_parity.py.

Per system, on a batch of harvested states (contacts active) -- and on the reference's golden
inputs -- one `System.step` of the CUDA path against the float32 oracle (o32) and the float64
oracle (o64 = the reference to 1e-11).  Per field:

  k-o32 max / p99   |kernel - o32|                      o32-o64 max / p99   |o32 - o64|
  k-o64 max / p99   |kernel - o64|                      (the float32 oracle's own error)
  literal           fraction of elements with |kernel - o32| <= 1e-6 + 1e-5 |o32|
  lit|acc           ... or kernel at least as close to o64 as o32 is (+ the literal tolerance)
and per system the envs outside the policy tolerance (`outliers`), as the tests count them.

usage: python tools/parity_report.py [--b 2048] [--steps 20] [names ...]   > profiles/rNN_parity_report.txt
The oracle is the checker here (tools/ is test infrastructure, like tests/).
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import util  # noqa: E402
from tests.test_gpu_parity import (ACCUMULATORS, GAP_MULT, VEL_ULPS, _env_gap, _flat, _harvest, _oracle,  # noqa: E402
                                   _system)

FIELDS = util.QP_KEYS + ('contact_vel', 'actuator_ang', 'contact_pos', 'contact_penetration')


def report(name, ins, act, got, tag):
  o32 = _oracle(name, np.float32).step(*ins, act)
  o64 = _oracle(name, np.float64).step(*ins, act)
  b = o32['pos'].shape[0]
  floor = VEL_ULPS * 6e-8 / util.load_desc(name).dt_sub
  bad = np.zeros(b, bool)
  print('== %s [%s]  B=%d  dt_sub=%g' % (name, tag, b, util.load_desc(name).dt_sub))
  for k in util.QP_KEYS + util.INFO_KEYS:
    if o32[k].size == 0:
      continue
    e32 = np.abs(got[k] - o32[k])
    e64 = np.abs(got[k] - o64[k])
    g = np.abs(o32[k] - o64[k])
    mag = np.abs(o32[k])
    if k in ACCUMULATORS:
      mag = np.broadcast_to(mag.reshape(b, -1).max(axis=1).reshape((b,) + (1,) * (mag.ndim - 1)), mag.shape)
    base = 1e-6 + 1e-5 * mag
    tol = base + GAP_MULT * _env_gap(o32[k], o64[k])
    if k not in ('pos', 'rot', 'contact_pos', 'contact_normal', 'contact_penetration'):
      tol = tol + floor
    bad |= (e32 > tol).reshape(b, -1).any(axis=1)
    if k in FIELDS:
      q = lambda x: (float(x.max()), float(np.quantile(x, 0.99)))  # noqa: E731
      print('  %-20s k-o32 %.1e / %.1e   k-o64 %.1e / %.1e   o32-o64 %.1e / %.1e   literal %.4f  lit|acc %.4f' % (
          (k,) + q(e32) + q(e64) + q(g) + (float((e32 <= base).mean()), float(((e32 <= base) | (e64 <= base + g)).mean()))))
  pen = o32['contact_penetration']
  if pen.size:
    mism = (got['contact_penetration'] > 0) != (pen > 0)
    hard = mism & (np.abs(pen) >= 1e-6) & (np.abs(o64['contact_penetration']) >= 1e-6)
    print('  contact flags: %d / %d differ, %d of them with |penetration| >= 1e-6' % (mism.sum(), mism.size, hard.sum()))
  print('  outliers: %d of %d envs outside tolerance (%.2f %%)' % (bad.sum(), b, 100.0 * bad.sum() / b))


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--b', type=int, default=2048)
  ap.add_argument('--steps', type=int, default=20)
  ap.add_argument('names', nargs='*')
  args = ap.parse_args()
  names = args.names or ['ant', 'humanoid', 'grasp', 'fetch', 'ant_spring', 'humanoid_spring', 'half_cheetah',
                         'humanoid_standup', 'pusher', 'ur5e', 'walker2d', 'hopper']
  print('lib', os.environ.get('BRAX_B200_LIB', 'default'))
  for name in names:
    s = _system(name)
    if name in util.golden_names():
      g = util.load_golden(name)
      ins = [g[k].astype(np.float32) for k in util.QP_KEYS]
      act = g['act'].astype(np.float32)
      from brax_b200.system import QP
      report(name, ins, act, _flat(*s.step(QP(*ins), act)), 'golden inputs')
    steps = 6 if name == 'half_cheetah' else args.steps
    qp, rng = _harvest(s, args.b, steps, 5)
    act = rng.uniform(-1, 1, (args.b, s.act_size)).astype(np.float32)
    got = _flat(*s.step(qp, torch.from_numpy(act).cuda()))
    ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
    report(name, ins, act, got, 'harvested, %d steps of random actions' % steps)


if __name__ == '__main__':
  main()
