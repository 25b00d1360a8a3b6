"""Adds tests/golden/reset_obs.npz: the obs the REFERENCE's `Env.reset` returns,
`_get_obs(qp, sys.info(qp))` (envs/ant.py:203-204, envs/humanoid.py:230-231 with a zero action,
envs/fetch.py:41-42, envs/grasp.py:60-61), evaluated by the unmodified env classes on the reset
states of tests/golden/reset.npz (numpy float64 through oracle/refsys.py), together with the
`info.contact` of `sys.info(qp)` those obs were built from.  Build container only.

usage: python oracle/gen_golden_reset_obs.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from gen_golden_env import VARIANTS, env_module  # noqa: E402


def main():
  g = np.load(os.path.join(ROOT, 'tests', 'golden', 'reset.npz'))
  rec = {}
  for name, (sysname, modname, clsname, kw) in VARIANTS.items():
    mod = env_module(modname)
    import brax
    e = getattr(mod, clsname)(**kw)
    obs, cv, ca = [], [], []
    for d in range(g[sysname + '/pos'].shape[0]):
      qp = brax.QP(*[g[sysname + '/' + k][d] for k in ('pos', 'rot', 'vel', 'ang')])
      info = e.sys.info(qp)
      if modname == 'humanoid':
        o = e._get_obs(qp, info, np.zeros(e.action_size))
      else:
        o = e._get_obs(qp, info)
      obs.append(np.asarray(o, np.float64))
      cv.append(np.asarray(info.contact.vel, np.float64))
      ca.append(np.asarray(info.contact.ang, np.float64))
    rec[name + '/obs'], rec[name + '/contact_vel'], rec[name + '/contact_ang'] = np.stack(obs), np.stack(cv), np.stack(ca)
    print('%-16s obs %s  max|info.contact| %.3g' % (name, rec[name + '/obs'].shape, max(np.abs(cv).max(), np.abs(ca).max())))
  np.savez_compressed(os.path.join(ROOT, 'tests', 'golden', 'reset_obs.npz'), **rec)


if __name__ == '__main__':
  main()
