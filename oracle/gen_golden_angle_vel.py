"""Adds tests/golden/<name>_angle_vel.npz: the REFERENCE's `Joint.angle_vel`
(brax/physics/joints.py:197-226, inherited by spring_joints.py) evaluated on the
OUTPUT states already stored in tests/golden/<name>_steps.npz (build container
only; unmodified /root/reference on numpy float64 through oracle/refsys.py).

Stored per system, in the flat joint order of the descriptors (reference joint
groups in `sys.joints` order, joints inside a group in config order):
  angle, vel  [S, nj, 3]  before the free_dofs selection (slots >= dof are 0)
  dof         [nj]        dofs of the joint's group
  free        [nj]        free_dofs of the joint (= dof when the group has none)
  flat_angle, flat_vel    list-concatenation of what `j.angle_vel(qp)` returns for
                          every group (what the envs' _get_obs consume), [S, total]

usage: python oracle/gen_golden_angle_vel.py [name ...]
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402
from gen_golden import SYSTEMS  # noqa: E402


def main(names):
  out_dir = os.path.join(ROOT, 'tests', 'golden')
  for name in names:
    env, var, _, _ = SYSTEMS[name]
    s = refsys.env_system(env, var)
    import brax
    g = np.load(os.path.join(out_dir, name + '_steps.npz'))
    steps = g['pos_o'].shape[0]
    dof, free = [], []
    for j in s.joints:
      n = len(j.index)
      dof += [j.dof] * n
      free += list(j.free_dofs) if j.free_dofs is not None else [j.dof] * n
    nj = len(dof)
    angle, vel = np.zeros((steps, nj, 3)), np.zeros((steps, nj, 3))
    flat_a, flat_v = [], []
    for t in range(steps):
      qp = brax.QP(pos=g['pos_o'][t], rot=g['rot_o'][t], vel=g['vel_o'][t], ang=g['ang_o'][t])
      fa, fv, base = [], [], 0
      for j in s.joints:
        a, v = j.angle_vel(qp)                    # what the envs consume
        fa.append(np.asarray(a, np.float64).reshape(-1))
        fv.append(np.asarray(v, np.float64).reshape(-1))
        keep, j.free_dofs = j.free_dofs, None     # the unselected values, per joint
        a, v = j.angle_vel(qp)
        j.free_dofs = keep
        n = len(j.index)
        angle[t, base:base + n, :j.dof] = np.asarray(a, np.float64).reshape(n, j.dof)
        vel[t, base:base + n, :j.dof] = np.asarray(v, np.float64).reshape(n, j.dof)
        base += n
      flat_a.append(np.concatenate(fa))
      flat_v.append(np.concatenate(fv))
    np.savez_compressed(os.path.join(out_dir, name + '_angle_vel.npz'), angle=angle, vel=vel,
                        dof=np.asarray(dof, np.int32), free=np.asarray(free, np.int32),
                        flat_angle=np.stack(flat_a), flat_vel=np.stack(flat_v))
    print('%-20s %3d states, %d joints, dof %s free %s' % (name, steps, nj, dof, free))


if __name__ == '__main__':
  main(sys.argv[1:] or list(SYSTEMS))
