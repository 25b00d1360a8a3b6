"""Generates tests/golden/friction/: reference steps under per-env collider friction (build container only).

The reference randomises friction per env with `DomainRandomizationWrapper(env_fn, *friction_randomizer(env,
values))` (/root/reference/brax/experimental/tracing/wrappers.py:27-88, randomizers.py:26-51): under jax.vmap every
env builds its System from `TracedConfig(config, custom_tree)` (tracing/customize.py:23-48) whose tree sets
material.friction of EVERY collider to that env's value.  jax.vmap is unavailable here, so each value builds
its own System through the reference's OWN TracedConfig + `post_process_fn` hook (physics/system.py:56-61) --
exactly what one vmap lane computes -- on the unmodified reference (numpy float64, oracle/refsys.py).

Per system (ant: pbd one-way capsule-plane; grasp: pbd two-way capsule-capsule + plane; ant_spring: legacy
spring) and friction value: from one common sliding start state (the default pose dropped onto the ground with a
horizontal velocity), T chained steps with fixed random actions; every step's input state and outputs (QP + Info)
are stored, plus the pair frictions `brax_b200.extract.describe` reads off the built System.

usage: python oracle/gen_golden_friction.py [system ...]
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
from gen_golden_bounce import _info_arrays  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden', 'friction')
FRICTION = (0.25, 0.6, 1.0, 1.7)
SYSTEMS = {'ant': ('ant', '_SYSTEM_CONFIG'), 'ant_spring': ('ant', '_SYSTEM_CONFIG_SPRING'),
           'grasp': ('grasp', '_SYSTEM_CONFIG')}
WARM, T = 6, 4


def friction_tree(config, f):
  """randomizers.py:39-47 with a scalar instead of the vmapped vector."""
  tree = {'bodies': []}
  for b in config.bodies:
    tree['bodies'].append({'colliders': [{'material': {'friction': f}} for _ in b.colliders]})
  return tree


def main(only):
  brax = refsys.setup()
  from brax.experimental.tracing.customize import TracedConfig
  from brax_b200.extract import describe
  os.makedirs(OUT, exist_ok=True)
  for name, (env, var) in SYSTEMS.items():
    if only and name not in only:
      continue
    t0 = time.time()
    config = refsys.parse(refsys.env_config_text(env, var))
    rng = np.random.default_rng(11)
    base = brax.System(config)
    qp0 = base.default_qp()
    nact = int(sum(np.asarray(a.act_index).size for a in base.actuators)) if base.actuators else 0
    nact = max(nact, 1)
    acts = rng.uniform(-1, 1, (WARM + T, 64))
    rec = {k: [] for k in ('pos', 'rot', 'vel', 'ang', 'act')}
    rec.update({'out_' + k: [] for k in ('pos', 'rot', 'vel', 'ang', 'contact_vel', 'contact_ang', 'joint_vel',
                                         'joint_ang', 'actuator_vel', 'actuator_ang', 'contact_pos', 'contact_normal',
                                         'contact_penetration')})
    pair_fr = []
    for f in FRICTION:
      s = brax.System(config, post_process_fn=lambda c, f=f: TracedConfig(c, friction_tree(c, f)))
      desc = describe(s, name, act_size=None) if 'act_size' in describe.__code__.co_varnames else describe(s, name)
      pair_fr.append([[p['friction_a'], p['friction_b']] for p in desc.pairs])
      a = desc.act_size
      qp = qp0
      # sliding start: free bodies get a horizontal velocity
      import dataclasses
      free = np.asarray([b['pos_mask'][0] for b in desc.bodies], np.float64)[:, None]
      qp = dataclasses.replace(qp, vel=np.asarray(qp.vel) + free * np.array([1.5, 0.5, 0.0])) if dataclasses.is_dataclass(qp) \
          else qp.__class__(pos=qp.pos, rot=qp.rot, vel=np.asarray(qp.vel) + free * np.array([1.5, 0.5, 0.0]), ang=qp.ang)
      per = {k: [] for k in rec}
      for t in range(WARM + T):
        act = acts[t, :a]
        if t >= WARM:
          for k in ('pos', 'rot', 'vel', 'ang'):
            per[k].append(np.asarray(getattr(qp, k), np.float64))
          per['act'].append(act.copy())
        qp, info = s.step(qp, act)
        if t >= WARM:
          for k in ('pos', 'rot', 'vel', 'ang'):
            per['out_' + k].append(np.asarray(getattr(qp, k), np.float64))
          for k, v in _info_arrays(info).items():
            per['out_' + k].append(v)
      for k in rec:
        rec[k].append(np.stack(per[k]))
    out = {k: np.stack(v) for k, v in rec.items()}      # [F, T, ...]
    out['friction'] = np.asarray(FRICTION)
    out['pair_friction'] = np.asarray(pair_fr)          # [F, P, 2] as extracted from the built Systems
    assert all(np.allclose(out['pair_friction'][i], np.float32(f)) for i, f in enumerate(FRICTION)), out['pair_friction']
    np.savez_compressed(os.path.join(OUT, name + '.npz'), **out)
    spread = np.abs(out['out_pos'][0, -1] - out['out_pos'][-1, -1]).max()
    active = (out['out_contact_penetration'] > 0).mean()
    print('%-12s %.0fs  F=%d T=%d  |pos(f=%.2f) - pos(f=%.2f)| after %d steps = %.3e  active contact rows %.0f %%' % (
        name, time.time() - t0, len(FRICTION), T, FRICTION[0], FRICTION[-1], WARM + T, spread, 100 * active))


JOINT_DELTAS = ((0.0, 0.0, 0.0), (0.02, -0.01, 0.015), (-0.03, 0.02, 0.0), (0.01, 0.03, -0.02))
JOINT_SYSTEMS = {'ant': ('ant', '_SYSTEM_CONFIG'), 'ant_spring': ('ant', '_SYSTEM_CONFIG_SPRING'),
                 'humanoid_spring': ('humanoid', '_SYSTEM_CONFIG_SPRING')}


def joint_tree(config, delta, randomized):
  """randomizers.py:73-93 with a scalar offset instead of the vmapped vector: parent_offset + delta on the
  selected joints, the config's own value on the others."""
  tree = {'joints': []}
  for i, j in enumerate(config.joints):
    d = delta if i in randomized else (0.0, 0.0, 0.0)
    tree['joints'].append({'parent_offset': {'x': d[0] + j.parent_offset.x, 'y': d[1] + j.parent_offset.y,
                                             'z': d[2] + j.parent_offset.z}})
  return tree


def main_joints(only):
  """tests/golden/friction/<sys>_joints.npz: the joint randomiser (randomizers.py:54-97) through the reference's
  own TracedConfig hook, one System per offset; every other joint is randomised.  Each variant starts from ITS OWN
  `default_qp()` (the reference builds the pose from the randomised sockets), given a sliding velocity."""
  brax = refsys.setup()
  from brax.experimental.tracing.customize import TracedConfig
  from brax_b200.extract import describe
  os.makedirs(OUT, exist_ok=True)
  for name, (env, var) in JOINT_SYSTEMS.items():
    if only and name not in only:
      continue
    t0 = time.time()
    config = refsys.parse(refsys.env_config_text(env, var))
    rng = np.random.default_rng(13)
    randomized = set(range(0, len(config.joints), 2))
    acts = rng.uniform(-1, 1, (WARM + T, 64))
    keys = ('pos', 'rot', 'vel', 'ang', 'act') + tuple('out_' + k for k in (
        'pos', 'rot', 'vel', 'ang', 'contact_vel', 'contact_ang', 'joint_vel', 'joint_ang', 'actuator_vel', 'actuator_ang',
        'contact_pos', 'contact_normal', 'contact_penetration'))
    rec = {k: [] for k in keys}
    offs, resets = [], []
    reset_rng = (rng.uniform(-0.3, 0.3, 64), rng.uniform(-0.5, 0.5, 64))
    for delta in JOINT_DELTAS:
      s = brax.System(config, post_process_fn=lambda c, d=delta: TracedConfig(c, joint_tree(c, d, randomized)))
      desc = describe(s, name)
      offs.append([j['off_p'] for j in desc.joints])
      a = desc.act_size
      # the reset pose of this variant: default_qp(joint_angle, joint_velocity) with fixed random joint states
      # (system.py:112-242 reads the randomised parent_offset, :199-206)
      nd = len(s.default_angle())
      ja = s.default_angle() + reset_rng[0][:nd]
      rq = s.default_qp(joint_angle=ja, joint_velocity=reset_rng[1][:nd])
      resets.append({k: np.asarray(getattr(rq, k), np.float64) for k in ('pos', 'rot', 'vel', 'ang')})
      reset_in = (np.asarray(ja, np.float64), np.asarray(reset_rng[1][:nd], np.float64))
      qp = s.default_qp()
      free = np.asarray([b['pos_mask'][0] for b in desc.bodies], np.float64)[:, None]
      qp = qp.__class__(pos=qp.pos, rot=qp.rot, vel=np.asarray(qp.vel) + free * np.array([1.0, 0.3, 0.0]), ang=qp.ang)
      per = {k: [] for k in keys}
      for t in range(WARM + T):
        act = acts[t, :a]
        if t >= WARM:
          for k in ('pos', 'rot', 'vel', 'ang'):
            per[k].append(np.asarray(getattr(qp, k), np.float64))
          per['act'].append(act.copy())
        qp, info = s.step(qp, act)
        if t >= WARM:
          for k in ('pos', 'rot', 'vel', 'ang'):
            per['out_' + k].append(np.asarray(getattr(qp, k), np.float64))
          for k, v in _info_arrays(info).items():
            per['out_' + k].append(v)
      for k in keys:
        rec[k].append(np.stack(per[k]))
    out = {k: np.stack(v) for k, v in rec.items()}      # [V, T, ...]
    out['joint_offset'] = np.asarray(offs)              # [V, J, 3] as extracted from the built Systems
    for k in ('pos', 'rot', 'vel', 'ang'):
      out['reset_' + k] = np.stack([r[k] for r in resets])      # [V, N, .] default_qp(joint_angle, joint_velocity)
    out['reset_joint_angle'], out['reset_joint_velocity'] = reset_in
    base = out['joint_offset'][0]
    # the descriptor lists the joints in the reference's joint-group order (by type), the config in file order:
    # match them by their (parent, child) bodies
    bidx = {b.name: i for i, b in enumerate(config.bodies)}
    cfg_of = {(bidx[j.parent], bidx[j.child]): i for i, j in enumerate(config.joints)}
    rand_desc = {k for k, j in enumerate(desc.joints) if cfg_of[(j['body_p'], j['body_c'])] in randomized}
    out['randomized'] = np.asarray(sorted(rand_desc))
    for v, d in enumerate(JOINT_DELTAS):
      for j in range(base.shape[0]):
        want = base[j] + (np.asarray(d) if j in rand_desc else 0.0)
        assert np.allclose(out["joint_offset"][v, j], want, atol=1e-7), (v, j, out["joint_offset"][v, j], want)
    np.savez_compressed(os.path.join(OUT, name + '_joints.npz'), **out)
    print('%-16s %.0fs  variants=%d T=%d  %d of %d joints randomised  |pos(v0) - pos(v1)| = %.3e' % (
        name, time.time() - t0, len(JOINT_DELTAS), T, len(randomized), base.shape[0],
        np.abs(out['out_pos'][0, -1] - out['out_pos'][1, -1]).max()))


if __name__ == '__main__':
  args = [a for a in sys.argv[1:] if a != '--joints']
  if '--joints' in sys.argv[1:]:
    main_joints(args)
  else:
    main(args)
