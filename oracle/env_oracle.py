"""TEST INFRASTRUCTURE ONLY: numpy restatement of what the reference's envs compute
AFTER `sys.step` inside `Env.step` -- obs, reward, done, metrics.

  ant       brax/envs/ant.py:222-281       (Ant.step :222-256, _get_obs :257-281)
  humanoid  brax/envs/humanoid.py:246-340  (Humanoid.step :246-279, _get_obs :281-330,
                                            _center_of_mass :332-334)
  fetch     brax/envs/fetch.py:54-122      (Fetch.step :54-96, _get_obs :98-122)
  grasp     brax/envs/grasp.py:73-186      (Grasp.step :73-138 incl. the action map and palm
                                            move BEFORE sys.step = env_prologue, _get_obs :144-186)
The random target teleport on a hit (fetch.py:88-93, grasp.py:131-136) draws from the JAX
PRNG and stays with the caller; `hits` is among the metrics.

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


def _rotate(v, q):
  """math.rotate (math.py:25-40), batched: v [...,3], q [...,4] (w,x,y,z)."""
  s, u = q[..., :1], q[..., 1:]
  r = 2 * (np.sum(u * v, axis=-1, keepdims=True) * u) + (s * s - np.sum(u * u, axis=-1, keepdims=True)) * v
  return r + 2 * s * np.cross(u, v)


def _inv_rotate(v, q):
  """math.inv_rotate (math.py:43-53): rotate by quat_inv(q) = q * [1,-1,-1,-1] (math.py:190-199)."""
  return _rotate(v, q * np.asarray([1, -1, -1, -1], q.dtype))


def _norm(x):
  return np.sqrt(np.sum(np.square(x), axis=-1))


def env_prologue(desc, env, pos, act, dtype=np.float32):
  """What `Env.step` does BEFORE `sys.step`: (pos, act) -> (pos', act').  Identity except
  for grasp (grasp.py:75-85): the [-1,1] action is mapped onto the joint ranges (+ the palm
  target box) and the frozen palm is moved towards its target, at most 0.15 * 2 per step."""
  dt = np.dtype(dtype).type
  pos, act = np.array(pos, dtype), np.asarray(act, dtype)
  if env.kind != 'grasp':
    return pos, act
  act = np.asarray(env.act_min, dtype) + np.asarray(env.act_range, dtype) * ((act + dt(1)) / dt(2.))
  target = act[:, -3:]
  palm = pos[:, env.ref_body]
  nrm = _norm(target - palm)
  scale = np.where(nrm > dt(2.0), dt(2.) / np.where(nrm > 0, nrm, dt(1)), dt(1.0)).astype(dtype)
  pos[:, env.ref_body] = palm + scale[:, None] * (target - palm) * dt(.15)
  return pos, act


def _egocentric(env, pos, rot, vel, cvel, dt):
  """Shared by fetch / grasp _get_obs: positions and velocities of all bodies in the frame
  of the reference body, contact flags (fetch.py:103-117, grasp.py:147-150,176-178)."""
  b = pos.shape[0]
  q = rot[:, env.ref_body][:, None, :]
  pos_local = _inv_rotate(pos - pos[:, env.ref_body][:, None, :], np.broadcast_to(q, rot.shape))
  vel_local = _inv_rotate(vel, np.broadcast_to(q, rot.shape))
  contacts = np.where(np.sum(np.square(cvel), axis=-1) > dt(0.00001), dt(1), dt(0)).astype(pos.dtype)
  return pos_local, vel_local, contacts


def env_teleport(env, pos, hit, rand, dtype=np.float32):
  """fetch.py:88-93 + :124-134, grasp.py:131-136 + :188-198: targets of envs with a hit move to
  (dist cos a, dist sin a, z), dist = radius + distance * u0, a = 2 pi u1, z = 1 | height * u2;
  rand [B,3] are the uniforms the reference draws from its PRNG key."""
  dt = np.dtype(dtype).type
  pos, rand = np.array(pos, dtype), np.asarray(rand, dtype)
  dist = dt(env.target_radius) + dt(env.target_distance) * rand[:, 0]
  ang = dt(np.pi) * dt(2.) * rand[:, 1]
  z = dt(env.target_height) * rand[:, 2] if env.target_height > 0 else np.ones_like(dist)
  target = np.stack([dist * np.cos(ang), dist * np.sin(ang), z], axis=-1).astype(dtype)
  pos[:, env.target_body] = np.where((np.asarray(hit) > 0)[:, None], target, pos[:, env.target_body])
  return pos


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
  elif env.kind == 'fetch':
    cvel = np.asarray(out['contact_vel'], dtype)
    ti, gi = env.ref_body, env.target_body
    pos_local, vel_local, contacts = _egocentric(env, pos, rot, vel, cvel, dt)
    ex = np.broadcast_to(np.asarray([1., 0., 0.], dtype), (b, 3))
    ez = np.broadcast_to(np.asarray([0., 0., 1.], dtype), (b, 3))
    torso_fwd, torso_up = _rotate(ex, rot[:, ti]), _rotate(ez, rot[:, ti])
    tl = pos_local[:, gi]
    tl_mag = _norm(tl)[:, None]
    obs = np.concatenate([torso_fwd, torso_up, tl_mag, tl / (dt(1e-6) + tl_mag), pos_local.reshape(b, -1),
                          vel_local.reshape(b, -1), contacts], axis=-1)
    torso_delta = pos[:, ti] - pos0[:, ti]                          # fetch.py:59-63
    target_rel = pos[:, gi] - pos[:, ti]
    target_dist = _norm(target_rel)
    target_dir = target_rel / (dt(1e-6) + target_dist)[:, None]
    moving = dt(.1) * np.sum(torso_delta * target_dir, axis=-1)
    is_up = dt(.1) * cdt * np.sum(torso_up * ez, axis=-1)
    height = dt(.1) * cdt * pos[:, 0, 2]
    facing = np.sum(target_dir * torso_fwd, axis=-1)
    hit = np.where(target_dist < dt(env.target_radius), dt(1), dt(0)).astype(dtype)
    weighted = hit * facing
    reward = height + moving + is_up + weighted                     # fetch.py:81
    done = np.zeros(b, dtype)
    metrics = np.stack([hit, weighted, moving, is_up, height], axis=-1)
  elif env.kind == 'grasp':
    cvel = np.asarray(out['contact_vel'], dtype)
    pi, oi, gi, hi = env.ref_body, env.object_body, env.target_body, env.hand_body
    pos_local, vel_local, contacts = _egocentric(env, pos, rot, vel, cvel, dt)

    def mag_dir(x):
      m = _norm(x)[:, None]
      return m, x / (dt(1e-6) + m)

    ol_mag, ol_dir = mag_dir(pos_local[:, oi])
    tl_mag, tl_dir = mag_dir(pos_local[:, gi])
    hand_to_obj = pos[:, oi] - pos[:, pi]
    _, h2o_dir = mag_dir(hand_to_obj)
    hand_vel = vel[:, hi]
    heading = np.sum(h2o_dir * hand_vel, axis=-1, keepdims=True)
    o2t_mag, o2t_dir = mag_dir(pos[:, gi] - pos[:, oi])
    obj_heading = np.sum(o2t_dir * vel[:, oi], axis=-1, keepdims=True)
    obs = np.concatenate([ol_mag, ol_dir, tl_mag, tl_dir, pos_local.reshape(b, -1), vel_local.reshape(b, -1),
                          hand_to_obj, hand_vel, heading, o2t_mag, o2t_dir, obj_heading, contacts], axis=-1)
    object_rel = pos[:, oi] - pos[:, pi]                             # grasp.py:91-100
    object_dist = _norm(object_rel)
    planar = _norm(object_rel * np.asarray([1., 1., 0.], dtype))
    object_dir = object_rel / (dt(1e-6) + object_dist)[:, None]
    moving_to_object = dt(.1) * cdt * np.sum(hand_vel * object_dir, axis=-1)
    close = dt(.1) * cdt * dt(1.) / (dt(1.) + planar)
    target_rel = pos[:, gi] - pos[:, oi]                             # grasp.py:103-109
    target_dist = _norm(target_rel)
    target_dir = target_rel / (dt(1e-6) + target_dist)[:, None]
    moving_to_target = dt(1.5) * cdt * np.sum(vel[:, oi] * target_dir, axis=-1)
    touching = dt(0.2) * cdt * np.sum(contacts[:, list(env.touch_bodies)], axis=-1)   # grasp.py:112-116
    hit = np.where(target_dist < dt(env.target_radius), dt(1), dt(0)).astype(dtype)
    reward = moving_to_object + close + touching + dt(5.) * hit + moving_to_target      # grasp.py:121
    done = np.zeros(b, dtype)
    metrics = np.stack([hit, touching, moving_to_object, moving_to_target, close], axis=-1)
  else:
    raise ValueError('no env epilogue for kind %r' % env.kind)
  return {'obs': obs.astype(dtype), 'reward': reward.astype(dtype), 'done': done.astype(dtype),
          'metrics': metrics.astype(dtype)}


def wrapped_step(desc, env, oracle, qp, done_in, steps_in, act, first_qp, first_obs, episode_length,
                 dtype=np.float32, rand=None, action_repeat=1):
  """One step of AutoResetWrapper(EpisodeWrapper(env, episode_length, action_repeat))
  (brax/envs/wrappers.py:131-148 around :105-120), as `envs.create` composes them
  (brax/envs/__init__.py:74-90).  qp / first_qp: dicts pos rot vel ang [B,N,*];
  oracle: the C physics oracle of the same dtype; rand: [B,3] or, with action_repeat > 1,
  [action_repeat,B,3] (one draw per env step).  Returns qp (dict), obs, reward, done, metrics,
  steps, truncation."""
  dt = np.dtype(dtype).type
  done_in, steps = np.asarray(done_in, dtype), np.asarray(steps_in, dtype)
  steps = np.where(done_in > 0, dt(0), steps)                       # wrappers.py:133-135
  cur = {k: np.asarray(qp[k], dtype) for k in ('pos', 'rot', 'vel', 'ang')}
  reward = None
  for rep in range(action_repeat):                                  # wrappers.py:106-110
    pos_in, act_in = env_prologue(desc, env, cur['pos'], act, dtype)
    out = oracle.step(pos_in, cur['rot'], cur['vel'], cur['ang'], act_in)
    fa, fv = flat_angle_vel(desc, *oracle.angle_vel(out['rot'], out['ang']))
    r = env_epilogue(desc, env, cur['pos'], out, act, fa, fv, dtype)
    if rand is not None:
      rr = np.asarray(rand)
      out['pos'] = env_teleport(env, out['pos'], r['metrics'][:, 0], rr[rep] if rr.ndim == 3 else rr, dtype)
    reward = r['reward'] if reward is None else (reward + r['reward']).astype(dtype)   # :111
    cur = {k: out[k] for k in ('pos', 'rot', 'vel', 'ang')}
  steps = steps + dt(action_repeat)                                 # wrappers.py:112
  over = steps >= dt(episode_length)
  done = np.where(over, dt(1), r['done']).astype(dtype)             # wrappers.py:116
  trunc = np.where(over, dt(1) - r['done'], dt(0)).astype(dtype)    # wrappers.py:117-118
  d = done > 0
  new_qp = {k: np.where(d.reshape((-1,) + (1,) * (out[k].ndim - 1)), np.asarray(first_qp[k], dtype), out[k])
            for k in ('pos', 'rot', 'vel', 'ang')}                  # wrappers.py:146
  obs = np.where(d[:, None], np.asarray(first_obs, dtype), r['obs'])  # wrappers.py:147
  return {'qp': new_qp, 'obs': obs.astype(dtype), 'reward': reward, 'done': done, 'metrics': r['metrics'],
          'steps': steps.astype(dtype), 'truncation': trunc}


def reset_obs(desc, env, oracle, qp, dtype=np.float32, contact_vel=None, contact_ang=None):
  """The obs of `Env.reset`: `_get_obs(qp, info)` of a state (envs/ant.py:203-204,
  envs/humanoid.py:230-231 with a zero action, envs/fetch.py:41-42, envs/grasp.py:60-61) with the
  given `info.contact` (default: that of the reference's `sys.info(qp)`, system.py:327-340 -- the
  spring-style contact impulse of the pose, `Oracle.info_contact`; zero unless a collider sits in
  the plane moving downwards).  oracle: the C physics oracle of the same dtype (Joint.angle_vel,
  info_contact)."""
  b, n = qp['pos'].shape[0], qp['pos'].shape[1]
  out = {k: np.asarray(qp[k], dtype) for k in ('pos', 'rot', 'vel', 'ang')}
  if contact_vel is None or contact_ang is None:
    ic = oracle.info_contact(out['pos'], out['rot'], out['vel'], out['ang'])   # sys.info(qp).contact
    contact_vel = ic['contact_vel'] if contact_vel is None else contact_vel
    contact_ang = ic['contact_ang'] if contact_ang is None else contact_ang
  out['contact_vel'] = np.asarray(contact_vel, dtype)
  out['contact_ang'] = np.asarray(contact_ang, dtype)
  fa, fv = flat_angle_vel(desc, *oracle.angle_vel(out['rot'], out['ang']))
  act = np.zeros((b, desc.act_size), dtype)
  if env.kind == 'grasp':     # the epilogue is given the raw action; zeros map to mid-range, unused by the obs
    act = np.zeros((b, desc.act_size), dtype)
  return env_epilogue(desc, env, out['pos'], out, act, fa, fv, dtype)['obs']


def eval_step(episode_metrics, active, episode_steps, metrics, reward, done, steps):
  """EvalWrapper.step's bookkeeping (brax/envs/wrappers.py:186-200) on arrays [B] / [B,M]:
  `metrics` are the step's env metrics with the reward appended last (`nstate.metrics['reward']`,
  :185).  Returns the new (episode_metrics [B,M+1], active_episodes, episode_steps)."""
  m = np.concatenate([np.asarray(metrics), np.asarray(reward)[:, None]], axis=-1)
  new_steps = np.where(active > 0, steps, episode_steps)            # :187-188
  new_metrics = episode_metrics + m * active[:, None]               # :189-191
  new_active = active * (1 - done)                                  # :192
  return new_metrics, new_active, new_steps
