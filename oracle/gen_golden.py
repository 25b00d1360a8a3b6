"""Generates tests/golden/*.npz by running the REFERENCE ITSELF (build container
only): the unmodified /root/reference brax.System.step on its numpy float64
backend (oracle/refsys.py).  These vectors pin the C oracle
(oracle/brax_step_oracle.c) and, through it, the CUDA path.

Per system: TRAJ rollouts from a noisy reset (the envs' own recipe,
envs/ant.py:198-204: default_angle + U(-.1,.1), velocity U(-.1,.1)) with
U(-1,1) actions; every step's input QP, action, output QP and Info is stored.

usage: python oracle/gen_golden.py [name ...]
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402

# name -> (env file, config variable, steps per trajectory, trajectories)
SYSTEMS = {
    'ant': ('ant', '_SYSTEM_CONFIG', 12, 2),
    'humanoid': ('humanoid', '_SYSTEM_CONFIG', 12, 2),
    'grasp': ('grasp', '_SYSTEM_CONFIG', 10, 2),
    'fetch': ('fetch', '_SYSTEM_CONFIG', 10, 2),
    'ant_spring': ('ant', '_SYSTEM_CONFIG_SPRING', 8, 2),
    'humanoid_spring': ('humanoid', '_SYSTEM_CONFIG_SPRING', 8, 2),
    'grasp_spring': ('grasp', '_SYSTEM_CONFIG_SPRING', 6, 1),
    'half_cheetah': ('half_cheetah', '_SYSTEM_CONFIG', 4, 1),
    'hopper': ('hopper', '_SYSTEM_CONFIG', 6, 1),
    'swimmer': ('swimmer', '_SYSTEM_CONFIG', 4, 1),
    'inverted_pendulum': ('inverted_pendulum', '_SYSTEM_CONFIG', 4, 1),
    'reacherangle': ('reacherangle', '_SYSTEM_CONFIG', 4, 1),
    'humanoid_standup': ('humanoid_standup', '_SYSTEM_CONFIG', 4, 1),
    # the remaining reference envs (every env of brax/envs with a v1 system is covered)
    'walker2d': ('walker2d', '_SYSTEM_CONFIG', 4, 1),
    'pusher': ('pusher', '_SYSTEM_CONFIG', 4, 1),
    'ur5e': ('ur5e', '_SYSTEM_CONFIG', 4, 1),
    'acrobot': ('acrobot', '_SYSTEM_CONFIG', 4, 1),
    'reacher': ('reacher', '_SYSTEM_CONFIG', 4, 1),
    'inverted_double_pendulum': ('inverted_double_pendulum', '_SYSTEM_CONFIG', 4, 1),
    'fetch_spring': ('fetch', '_SYSTEM_CONFIG_SPRING', 4, 1),
    'walker2d_spring': ('walker2d', '_SYSTEM_CONFIG_SPRING', 4, 1),
    'ur5e_spring': ('ur5e', '_SYSTEM_CONFIG_SPRING', 4, 1),
}
EXTRA_ACT = {'grasp': 3}


def main(names):
  out_dir = os.path.join(ROOT, 'tests', 'golden')
  os.makedirs(out_dir, exist_ok=True)
  for name in names:
    env, var, steps, trajs = SYSTEMS[name]
    t0 = time.time()
    s = refsys.env_system(env, var)
    a = int(s.num_joint_dof + s.num_forces_dof) + EXTRA_ACT.get(env, 0)
    nd = s.num_joint_dof
    rec = {k: [] for k in (
        'pos', 'rot', 'vel', 'ang', 'act', 'pos_o', 'rot_o', 'vel_o', 'ang_o',
        'contact_vel', 'contact_ang', 'joint_vel', 'joint_ang',
        'actuator_vel', 'actuator_ang', 'contact_pos', 'contact_normal',
        'contact_penetration')}
    for tr in range(trajs):
      rng = np.random.default_rng(1000 + tr)
      qp = s.default_qp(
          joint_angle=s.default_angle() + rng.uniform(-.1, .1, nd),
          joint_velocity=rng.uniform(-.1, .1, nd))
      for _ in range(steps):
        act = rng.uniform(-1, 1, a)
        qp2, info = s.step(qp, act)
        for k in ('pos', 'rot', 'vel', 'ang'):
          rec[k].append(np.asarray(getattr(qp, k), np.float64))
          rec[k + '_o'].append(np.asarray(getattr(qp2, k), np.float64))
        rec['act'].append(act)
        for k in ('contact', 'joint', 'actuator'):
          p = getattr(info, k)
          rec[k + '_vel'].append(np.asarray(p.vel, np.float64))
          rec[k + '_ang'].append(np.asarray(p.ang, np.float64))
        kk = np.asarray(info.contact_penetration).shape[0]
        rec['contact_pos'].append(
            np.asarray(info.contact_pos, np.float64).reshape(kk, 3))
        rec['contact_normal'].append(
            np.asarray(info.contact_normal, np.float64).reshape(kk, 3))
        rec['contact_penetration'].append(
            np.asarray(info.contact_penetration, np.float64).reshape(kk))
        qp = qp2
    np.savez_compressed(os.path.join(out_dir, name + '_steps.npz'),
                        **{k: np.stack(v) for k, v in rec.items()})
    print('%-20s %3d steps  %.1fs' % (name, steps * trajs, time.time() - t0))


if __name__ == '__main__':
  main(sys.argv[1:] or list(SYSTEMS))
