"""Prints max |kernel - oracle| per output for every golden system (GPU box).

usage: python tools/parity_report.py [--threads T]
The oracle is the checker here (tools/ is test infrastructure, like tests/).
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from brax_b200.system import QP, System  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from tests import util  # noqa: E402

KEYS = (('pos', 'pos_o'), ('rot', 'rot_o'), ('vel', 'vel_o'), ('ang', 'ang_o'))


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--threads', type=int, default=0)
  ap.add_argument('names', nargs='*')
  args = ap.parse_args()
  for name in (args.names or util.golden_names()):
    g = util.load_golden(name)
    d = util.load_desc(name)
    s = System(d)
    if args.threads:
      s.configure(args.threads)
    qp = QP(*[g[k].astype(np.float32) for k in ('pos', 'rot', 'vel', 'ang')])
    act = g['act'].astype(np.float32)
    qo, info = s.step(qp, act)
    o32 = Oracle(d, np.float32).step(qp.pos, qp.rot, qp.vel, qp.ang, act)
    o64 = Oracle(d, np.float64).step(qp.pos, qp.rot, qp.vel, qp.ang, act)
    got = {'pos': qo.pos, 'rot': qo.rot, 'vel': qo.vel, 'ang': qo.ang,
           'contact_vel': info.contact.vel, 'contact_ang': info.contact.ang,
           'joint_vel': info.joint.vel, 'joint_ang': info.joint.ang,
           'actuator_vel': info.actuator.vel, 'actuator_ang': info.actuator.ang,
           'contact_pos': info.contact_pos, 'contact_normal': info.contact_normal,
           'contact_penetration': info.contact_penetration}
    parts = []
    for k, v in got.items():
      if v.size == 0:
        continue
      e32 = np.abs(v - o32[k]).max()
      e64 = np.abs(v - o64[k]).max()
      o3264 = np.abs(o32[k] - o64[k]).max()
      parts.append('%s k-o32=%.1e k-o64=%.1e o32-o64=%.1e' % (k, e32, e64, o3264))
    flags = (got['contact_penetration'] > 0) != (o32['contact_penetration'] > 0)
    print('== %s finite=%s flag_mismatch=%d/%d' % (
        name, all(np.isfinite(v).all() for v in got.values()), int(flags.sum()), flags.size))
    for p in parts:
      print('   ' + p)


if __name__ == '__main__':
  main()
