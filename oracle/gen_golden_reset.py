"""Writes tests/golden/reset.npz: the REFERENCE's `System.default_qp(joint_angle, joint_velocity)`
(brax/physics/system.py:112-242, unmodified, numpy float64 through oracle/refsys.py) for every env
config of brax/envs, on DRAWS random joint states each: `default_angle() + U(-.4,.4)`, velocity
U(-.5,.5) -- the envs' own reset recipe (envs/ant.py:198-204) with wider noise.  Build container only.

usage: python oracle/gen_golden_reset.py
"""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402

DRAWS = 6


def main():
  rec = {}
  for f in sorted(glob.glob(os.path.join(ROOT, 'brax_b200', 'systems', '*.json'))):
    name = os.path.basename(f)[:-5]
    env, var = (name[:-7], '_SYSTEM_CONFIG_SPRING') if name.endswith('_spring') else (name, '_SYSTEM_CONFIG')
    s = refsys.env_system(env, var)
    nd = len(np.asarray(s.default_angle()))   # dofs by the config's angle limits (system.py:91-93)
    rng = np.random.default_rng(sum(map(ord, name)))
    out = {k: [] for k in ('angle', 'velocity', 'pos', 'rot', 'vel', 'ang')}
    for d in range(DRAWS):
      ja = np.asarray(s.default_angle(), np.float64) + (rng.uniform(-.4, .4, nd) if d else 0.0)
      jv = rng.uniform(-.5, .5, nd) if d else np.zeros(nd)
      qp = s.default_qp(joint_angle=ja, joint_velocity=jv)
      out['angle'].append(ja); out['velocity'].append(jv)
      for k in ('pos', 'rot', 'vel', 'ang'):
        out[k].append(np.asarray(getattr(qp, k), np.float64))
    for k, v in out.items():
      rec[name + '/' + k] = np.stack(v)
    print('%-32s D=%2d  z range [%.3f, %.3f]' % (name, nd, rec[name + '/pos'][..., 2].min(), rec[name + '/pos'][..., 2].max()))
  np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'reset.npz'), **rec)


if __name__ == '__main__':
  main()
