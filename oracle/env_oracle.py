"""TEST INFRASTRUCTURE ONLY: numpy restatement of what the reference's envs compute
AFTER `sys.step` inside `Env.step` -- obs, reward, done, metrics.

  ant       brax/envs/ant.py:222-281       (Ant.step :222-256, _get_obs :257-281)
  humanoid  brax/envs/humanoid.py:246-340  (Humanoid.step :246-279, _get_obs :281-330,
                                            _center_of_mass :332-334)

Pinned to the reference: tests/test_oracle_golden.py feeds it the physics outputs
stored in tests/golden/<sys>_steps.npz / <sys>_angle_vel.npz and compares with
tests/golden/<variant>_env.npz (the reference's own Env.step on numpy float64,
oracle/gen_golden_env.py).  Its float32 evaluation is the checker of the CUDA env
epilogue.  Only tests/, smoke() and bench.py's CPU legs may import this module.

    env_epilogue(desc, env, pos_before, out, act, flat_angle, flat_vel, dtype)
        desc        SystemDesc
        env         brax_b200.envs.EnvDesc (kind + the constructor arguments)
        pos_before  [B,N,3] qp.pos BEFORE the step
        out         dict of the step's outputs (oracle.step keys: pos rot vel ang contact_vel contact_ang)
        act         [B,A]
        flat_angle, flat_vel  [B,D] concatenation of `j.angle_vel(qp)` over joint groups
    -> dict(obs [B,O], reward [B], done [B], metrics [B,M])
"""
import numpy as np


def flat_angle_vel(desc, angle, vel):
  """[B,J,3] joint angles / velocities (Oracle.angle_vel) -> the concatenation the envs
  consume: per joint its first free_dofs dofs (joints.py:216-225)."""
  cols = [(i, d) for i, j in enumerate(desc.joints)
          for d in range(int(j.get('free_dofs', j['dof'])))]
  ia = np.asarray([c[0] for c in cols], int)
  da = np.asarray([c[1] for c in cols], int)
  return angle[:, ia, da], vel[:, ia, da]


def _healthy(env, z, dt):
  # ant.py:229-231, humanoid.py:255-257
  h = np.where(z < dt(env.healthy_z_range[0]), dt(0.0), dt(1.0))
  return np.where(z > dt(env.healthy_z_range[1]), dt(0.0), h).astype(dt)


def env_epilogue(desc, env, pos_before, out, act, flat_angle, flat_vel, dtype=np.float32):
  dt = np.dtype(dtype).type
  pos0 = np.asarray(pos_before, dtype)
  pos, rot = np.asarray(out['pos'], dtype), np.asarray(out['rot'], dtype)
  vel, ang = np.asarray(out['vel'], dtype), np.asarray(out['ang'], dtype)
  act = np.asarray(act, dtype)
  ja, jv = np.asarray(flat_angle, dtype), np.asarray(flat_vel, dtype)
  b = pos.shape[0]
  cdt = dt(desc.dt)
  if env.kind == 'ant':
    cvel = np.clip(np.asarray(out['contact_vel'], dtype), dt(-1), dt(1))
    cang = np.clip(np.asarray(out['contact_ang'], dtype), dt(-1), dt(1))
    velocity = (pos[:, 0] - pos0[:, 0]) / cdt                      # ant.py:226
    forward = velocity[:, 0]
    healthy = _healthy(env, pos[:, 0, 2], dt)
    hr = np.full(b, dt(env.healthy_reward)) if env.terminate_when_unhealthy else dt(env.healthy_reward) * healthy
    ctrl = dt(env.ctrl_cost_weight) * np.sum(np.square(act), axis=-1)
    contact = dt(env.contact_cost_weight) * np.sum(np.square(cvel), axis=(-1, -2))
    reward = forward + hr - ctrl - contact                          # ant.py:240
    done = (dt(1.0) - healthy) if env.terminate_when_unhealthy else np.zeros(b, dtype)
    qpos = [pos[:, 0, 2:] if env.exclude_current_positions else pos[:, 0], rot[:, 0], ja]
    qvel = [vel[:, 0], ang[:, 0], jv]
    cfrc = [cvel.reshape(b, -1), cang.reshape(b, -1)] if env.use_contact_forces else []
    obs = np.concatenate(qpos + qvel + cfrc, axis=-1)
    metrics = np.stack([forward, hr, -ctrl, -contact, pos[:, 0, 0], pos[:, 0, 1],
                        np.sqrt(np.sum(np.square(pos[:, 0]), axis=-1)),
                        velocity[:, 0], velocity[:, 1], forward], axis=-1)
  elif env.kind == 'humanoid':
    mass = np.asarray([x['mass'] for x in desc.bodies], dtype)
    inert = np.asarray([x['inv_inertia'] for x in desc.bodies], dtype)  # Body.inertia IS the inverse (bodies.py:40)
    msum = np.sum(mass[:-1])

    def com_of(p):                                                   # humanoid.py:332-334
      return np.sum(mass[:-1, None] * p[:, :-1], axis=1) / msum

    com_b, com_a = com_of(pos0), com_of(pos)
    velocity = (com_a - com_b) / cdt                                 # humanoid.py:252
    forward = dt(env.forward_reward_weight) * velocity[:, 0]
    healthy = _healthy(env, pos[:, 0, 2], dt)
    hr = np.full(b, dt(env.healthy_reward)) if env.terminate_when_unhealthy else dt(env.healthy_reward) * healthy
    ctrl = dt(env.ctrl_cost_weight) * np.sum(np.square(act), axis=-1)
    reward = forward + hr - ctrl
    done = (dt(1.0) - healthy) if env.terminate_when_unhealthy else np.zeros(b, dtype)
    # com_vals, humanoid.py:301-308 (norm, then squared)
    d = pos - com_a[:, None]
    n2 = np.sqrt(np.sum(np.square(d), axis=-1)) ** 2
    eye = np.eye(3, dtype=dtype)
    inr = (mass[None, :, None, None] * eye) * n2[..., None, None]
    inr = inr + inert[None, :, :, None] * eye - d[..., :, None] * d[..., None, :]
    com_vel = mass[None, :, None] * vel / msum
    com_ang = np.cross(d, vel) / (dt(1e-7) + n2)[..., None]
    qfrc = []
    groups = sorted(set(a['group'] for a in desc.actuators))
    for g in groups:                                                 # humanoid.py:314-319
      ga = [a for a in desc.actuators if a['group'] == g]
      idx = np.clip(np.asarray([a['act_index'] for a in ga], int), 0, act.shape[-1] - 1)  # jp.take, mode='clip'
      strength = np.asarray([a['strength'] for a in ga], dtype)
      t = act[:, idx.reshape(-1)] * np.repeat(strength, idx.shape[-1])
      qfrc.append(t.astype(dtype))
    qpos = [pos[:, 0, 2:] if env.exclude_current_positions else pos[:, 0], rot[:, 0], ja]
    qvel = [vel[:, 0], ang[:, 0], jv]
    obs = np.concatenate(qpos + qvel + [inr[:, :-1].reshape(b, -1), com_vel[:, :-1].reshape(b, -1),
                                        com_ang[:, :-1].reshape(b, -1)] + qfrc, axis=-1)
    metrics = np.stack([forward, forward, -ctrl, hr, com_a[:, 0], com_a[:, 1],
                        np.sqrt(np.sum(np.square(com_a), axis=-1)), velocity[:, 0], velocity[:, 1]], axis=-1)
  else:
    raise ValueError('no env epilogue for kind %r' % env.kind)
  return {'obs': obs.astype(dtype), 'reward': reward.astype(dtype), 'done': done.astype(dtype),
          'metrics': metrics.astype(dtype)}
