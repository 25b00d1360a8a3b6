"""Regenerates brax_b200/systems/*.json from the reference (build container only).

Runs the UNMODIFIED reference `brax.System` constructor on numpy (via
oracle/refsys.py) for every env config shipped in /root/reference/brax/envs and
serialises `brax_b200.extract.describe(sys)`.  The JSON files are committed so
that tests, bench.py and smoke() on the GPU box (no /root/reference there) can
create the same systems.

usage: python tools/gen_systems.py
"""
import glob
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))

import refsys  # noqa: E402
from brax_b200.extract import describe  # noqa: E402

# grasp's env appends 3 palm-translation slots to the action (grasp.py:138-140)
EXTRA_ACT = {'grasp': 3}


def main():
  out = os.path.join(ROOT, 'brax_b200', 'systems')
  os.makedirs(out, exist_ok=True)
  for f in sorted(glob.glob(os.path.join(refsys.REFERENCE, 'brax/envs/*.py'))):
    env = os.path.basename(f)[:-3]
    if env == 'humanoid_new':   # same physics config as humanoid
      continue
    for var, suffix in (('_SYSTEM_CONFIG', ''),
                        ('_SYSTEM_CONFIG_SPRING', '_spring')):
      try:
        refsys.env_config_text(env, var)
      except KeyError:
        continue
      s = refsys.env_system(env, var)
      a = int(s.num_joint_dof + s.num_forces_dof) + (
          EXTRA_ACT.get(env, 0))
      d = describe(s, env + suffix, act_size=a)
      d.validate()
      with open(os.path.join(out, env + suffix + '.json'), 'w') as fh:
        fh.write(d.to_json())
      print('%-28s N=%2d J=%2d A=%2d K=%3d %s' % (
          env + suffix, d.num_bodies, d.num_joints, d.act_size,
          d.num_contacts, d.dynamics_mode))


if __name__ == '__main__':
  main()
