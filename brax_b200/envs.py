"""Host-side mirror of the reference's `brax.envs` for the envs whose whole
`Env.step` runs as the fused kernel's epilogue ("next" row N1 of SURVEY.md 8(f)).

    env = envs.Ant(use_contact_forces=True)        # brax/envs/ant.py:173-196
    env = envs.Humanoid()                          # brax/envs/humanoid.py:200-221
    state = env.reset(batch=8192, device='cuda')   # noisy joint state -> default_qp (FK) -> obs: two launches
    state = env.step(state, action)                # ONE launch: physics + obs/reward/done/metrics

`State` carries the reference's fields (brax/envs/env.py:28-37): qp, obs, reward,
done, metrics (dict, the reference's keys in the reference's order), info.
`env.sys` is the `brax_b200.System` (so `env.sys.step` stays available).

What runs on the GPU: `sys.step` and, in the same launch, everything the
reference computes after it in `Ant.step` / `Humanoid.step` / `Fetch.step` / `Grasp.step`
(ant.py:222-281, humanoid.py:246-340, fetch.py:54-122, grasp.py:73-186), optionally with
EpisodeWrapper + AutoResetWrapper folded in (`EpisodeAutoReset`); `reset` (row N3): the forward
kinematics of `default_qp(joint_angle, joint_velocity)` (system.py:112-242, bx_default_qp) and
the reset obs (`System.observe`).  Agents are callers and stay reference code.  No CPU fallback.
"""
import ctypes as C
import dataclasses
from typing import Any, Dict, Tuple

import numpy as np

from brax_b200 import system as _system
from brax_b200.desc import MAX_ACTUATORS, MAX_JOINTS, SystemDesc

ENV_KIND = {'ant': 1, 'humanoid': 2, 'fetch': 3, 'grasp': 4}

ANT_METRICS = ('reward_forward', 'reward_survive', 'reward_ctrl', 'reward_contact', 'x_position',
               'y_position', 'distance_from_origin', 'x_velocity', 'y_velocity', 'forward_reward')
HUMANOID_METRICS = ('forward_reward', 'reward_linvel', 'reward_quadctrl', 'reward_alive', 'x_position',
                    'y_position', 'distance_from_origin', 'x_velocity', 'y_velocity')
FETCH_METRICS = ('hits', 'weightedHits', 'movingToTarget', 'torsoIsUp', 'torsoHeight')
GRASP_METRICS = ('hits', 'touchingObject', 'movingToObject', 'movingObjectToTarget', 'closeToObject')
MAX_ACT = 64


class CEnvDesc(C.Structure):
  """bx_env_desc (include/brax_b200.h)."""
  _fields_ = [('struct_size', C.c_uint32), ('kind', C.c_int32), ('dt', C.c_float),
              ('forward_reward_weight', C.c_float), ('ctrl_cost_weight', C.c_float),
              ('contact_cost_weight', C.c_float), ('healthy_reward', C.c_float),
              ('healthy_z_min', C.c_float), ('healthy_z_max', C.c_float),
              ('terminate_when_unhealthy', C.c_int32), ('use_contact_forces', C.c_int32),
              ('exclude_current_positions', C.c_int32),
              ('joint_free_dofs', C.c_int32 * MAX_JOINTS),
              ('actuator_width', C.c_int32 * MAX_ACTUATORS),
              ('ref_body', C.c_int32), ('target_body', C.c_int32), ('object_body', C.c_int32),
              ('hand_body', C.c_int32), ('touch_bodies', C.c_int32 * 4), ('target_radius', C.c_float),
              ('target_distance', C.c_float), ('target_height', C.c_float),
              ('act_min', C.c_float * MAX_ACT), ('act_range', C.c_float * MAX_ACT)]


@dataclasses.dataclass
class EnvDesc:
  """Constructor arguments of a reference env that enter `Env.step`."""
  kind: str
  forward_reward_weight: float = 1.0
  ctrl_cost_weight: float = 0.5
  contact_cost_weight: float = 0.0
  healthy_reward: float = 1.0
  terminate_when_unhealthy: bool = True
  healthy_z_range: Tuple[float, float] = (0.2, 1.0)
  use_contact_forces: bool = False
  exclude_current_positions: bool = True
  # fetch / grasp (fetch.py:28-34, grasp.py:32-54): body indices, target radius and, for
  # grasp, the affine map of the [-1,1] action (`_min_act`, `_range_act`)
  ref_body: int = 0         # torso (fetch) / palm (grasp): origin of the egocentric frame
  target_body: int = 0
  object_body: int = 0
  hand_body: int = 0
  touch_bodies: Tuple[int, ...] = (0, 0, 0, 0)
  target_radius: float = 0.0
  target_distance: float = 0.0   # _random_target: dist = radius + distance * u0
  target_height: float = 0.0     # z = height * u2; <= 0: z = 1 (fetch)
  act_min: Tuple[float, ...] = ()
  act_range: Tuple[float, ...] = ()

  def metric_names(self):
    return {'ant': ANT_METRICS, 'humanoid': HUMANOID_METRICS, 'fetch': FETCH_METRICS,
            'grasp': GRASP_METRICS}[self.kind]

  def obs_size(self, desc: SystemDesc) -> int:
    if self.kind == 'fetch':
      return 10 + 7 * desc.num_bodies
    if self.kind == 'grasp':
      return 20 + 7 * desc.num_bodies
    nd = sum(int(j.get('free_dofs', j['dof'])) for j in desc.joints)
    n = (1 if self.exclude_current_positions else 3) + 4 + nd + 6 + nd
    if self.kind == 'ant':
      return n + (6 * desc.num_bodies if self.use_contact_forces else 0)
    return n + 15 * (desc.num_bodies - 1) + sum(len(a['act_index']) for a in desc.actuators)

  def to_c(self, desc: SystemDesc) -> CEnvDesc:
    c = CEnvDesc()
    c.struct_size = C.sizeof(CEnvDesc)
    c.kind = ENV_KIND[self.kind]
    c.dt = desc.dt
    c.forward_reward_weight = self.forward_reward_weight
    c.ctrl_cost_weight = self.ctrl_cost_weight
    c.contact_cost_weight = self.contact_cost_weight
    c.healthy_reward = self.healthy_reward
    c.healthy_z_min, c.healthy_z_max = self.healthy_z_range
    c.terminate_when_unhealthy = int(self.terminate_when_unhealthy)
    c.use_contact_forces = int(self.use_contact_forces)
    c.exclude_current_positions = int(self.exclude_current_positions)
    for i, j in enumerate(desc.joints):
      c.joint_free_dofs[i] = int(j.get('free_dofs', j['dof']))
    for i, a in enumerate(desc.actuators):
      c.actuator_width[i] = len(a['act_index'])
    c.ref_body, c.target_body = self.ref_body, self.target_body
    c.object_body, c.hand_body = self.object_body, self.hand_body
    c.touch_bodies[:] = list(self.touch_bodies)
    c.target_radius = self.target_radius
    c.target_distance, c.target_height = self.target_distance, self.target_height
    if self.kind == 'grasp':
      if len(self.act_min) != desc.act_size or len(self.act_range) != desc.act_size:
        raise ValueError('grasp: act_min / act_range must have act_size entries')
      for i in range(desc.act_size):
        c.act_min[i], c.act_range[i] = self.act_min[i], self.act_range[i]
    return c


def fetch_desc(desc: SystemDesc) -> EnvDesc:
  """EnvDesc of `Fetch` (fetch.py:28-34): torso / target indices by body name, radius 2."""
  names = desc.body_names
  return EnvDesc('fetch', ref_body=names.index('Torso'), target_body=names.index('Target'), target_radius=2.0,
                 target_distance=15.0)


def grasp_desc(desc: SystemDesc) -> EnvDesc:
  """EnvDesc of `Grasp` (grasp.py:32-54).  The action map: per joint dof the angle limits in
  DEGREES as they stand in `sys.config.joints` after `joints.get` sphericalized them (three per
  joint, zeros appended, joints.py:431-433), then the palm target box."""
  names = desc.body_names
  if desc.dynamics_mode != 'pbd':
    # legacy_spring keeps 1/2/3-dof joints: the env lists the limits in CONFIG joint order, the
    # descriptor holds the joints grouped by dof (joints.py:465-474) -- not recoverable here
    raise NotImplementedError('grasp env epilogue: pbd only')
  lo, rng = [], []
  for j in desc.joints:
    lim = np.asarray(j['limit'], np.float64).reshape(-1, 2) * 180.0 / np.pi
    lo += [float(x) for x in lim[:, 0]]
    rng += [float(x) for x in lim[:, 1] - lim[:, 0]]
  lo += [-10.0, -10.0, 3.5]
  rng += [20.0, 20.0, 10.0]
  return EnvDesc('grasp', ref_body=names.index('HandPalm'), target_body=names.index('Target'),
                 object_body=names.index('Object'), hand_body=names.index('HandThumbProximal'),
                 touch_bodies=(3, 9, 12, 15),   # grasp.py:116, as written there
                 target_radius=1.1, target_distance=10.0, target_height=8.0, act_min=tuple(lo), act_range=tuple(rng))


@dataclasses.dataclass
class State:
  """Environment state (brax/envs/env.py:28-37)."""
  qp: Any
  obs: Any
  reward: Any
  done: Any
  metrics: Dict[str, Any] = dataclasses.field(default_factory=dict)
  info: Dict[str, Any] = dataclasses.field(default_factory=dict)

  def replace(self, **kw):
    return dataclasses.replace(self, **kw)


class Env:
  """Common part of the mirrored envs: a `System` with the env epilogue switched on."""

  def __init__(self, system_name: str, env, device: int = 0):
    """env: an EnvDesc, or a function SystemDesc -> EnvDesc (fetch_desc, grasp_desc)."""
    import os
    if callable(env):
      path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'systems', system_name + '.json')
      env = env(SystemDesc.load(path))
    self.env = env
    # True: state.info['physics'] carries the step's Info (an extra of this mirror; the reference's
    # Env.step drops it, ant.py:224) -- costs the Info stores of the launch
    self.keep_info = False
    # per-env collider friction [B] (CUDA tensor) or None: see `randomize_friction`
    self.friction = None
    # per-env joint sockets [B,J,3] (CUDA tensor) or None: see `randomize_joints`
    self.joint_offset = None
    self.sys = _system.load_system(system_name, device, env)
    self.observation_size = env.obs_size(self.sys.desc)
    self.action_size = self.sys.act_size

  def reset(self, batch: int = None, device=None, rng=None, velocity_noise: float = 0.0, seed: int = None,
            joint_angle=None, joint_velocity=None) -> State:
    """`Env.reset` (ant.py:198-220, humanoid.py:223-245, fetch.py:36-53, grasp.py:55-72).

    On a CUDA `device` this is the reference's recipe in two launches: joint angles
    `default_angle() + U(-s, s)` and joint velocities `U(-s, s)` per env (s = reset_noise_scale;
    torch generator seeded with `seed` -- same distribution as the reference's JAX draws, not the
    same stream; pass `joint_angle=` / `joint_velocity=` [B,D] to supply your own) ->
    `sys.default_qp(joint_angle, joint_velocity)` (bx_default_qp: forward kinematics) ->
    fetch: a random target (fetch.py:38-40) -> obs = `_get_obs(qp, info)` with a zero Info
    (`System.observe`).  reward = done = 0 and zero metrics as in the reference.

    Without a device (numpy): the recorded `default_qp()` with optional noise on the body
    velocities only, obs zeros (the host path has no reset kernel)."""
    if device is not None and str(device).startswith('cuda') and self.sys.desc.reset:
      return self._reset_cuda(batch, device, seed, joint_angle, joint_velocity)
    qp = self.sys.default_qp(batch=batch)
    if velocity_noise:
      rng = rng or np.random.default_rng(0)
      d = self.sys.desc
      mp = np.asarray([x['pos_mask'] for x in d.bodies], np.float32)
      mr = np.asarray([x['rot_mask'] for x in d.bodies], np.float32)
      qp.vel = (rng.uniform(-velocity_noise, velocity_noise, qp.pos.shape).astype(np.float32) * mp)
      qp.ang = (rng.uniform(-velocity_noise, velocity_noise, qp.pos.shape).astype(np.float32) * mr)
    lead = qp.pos.shape[:-2]
    zeros = lambda *s: np.zeros(lead + s, np.float32)  # noqa: E731
    obs, rew, done = zeros(self.observation_size), zeros(), zeros()
    met = {k: zeros() for k in self.env.metric_names()}
    if device is not None:
      import torch
      to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)  # noqa: E731
      qp = _system.QP(*[to(getattr(qp, k)) for k in ('pos', 'rot', 'vel', 'ang')])
      obs, rew, done = to(obs), to(rew), to(done)
      met = {k: to(v) for k, v in met.items()}
    return State(qp, obs, rew, done, met)

  def _reset_cuda(self, batch, device, seed, joint_angle, joint_velocity) -> State:
    import torch
    b = int(batch) if batch is not None else 1
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    if seed is None:
      # no seed given: every call draws fresh noise (a per-env counter on top of a fixed base), the way
      # the reference's callers split a new key per reset; pass `seed` for a reproducible state
      self._reset_count = getattr(self, '_reset_count', 0) + 1
      seed = 0x5EED0000 + self._reset_count
    gen.manual_seed(int(seed))
    nd, scale = self.sys.num_joint_dof, float(getattr(self, '_reset_noise_scale', 0.0))
    noise = lambda: (torch.rand(b, nd, generator=gen, device=dev, dtype=torch.float32) * 2 - 1) * scale  # noqa: E731
    if joint_angle is None:
      joint_angle = torch.from_numpy(self.sys.default_angle()).to(dev) + noise()    # ant.py:202
    if joint_velocity is None:
      joint_velocity = noise()                                                      # ant.py:203
    qp = self.sys.default_qp(joint_angle=joint_angle.reshape(b, nd), joint_velocity=joint_velocity.reshape(b, nd),
                             joint_offset=self.joint_offset)
    qp = self._place_targets(qp, gen)
    obs = self.sys.observe(qp, self.friction)
    names = self.env.metric_names()
    zeros = torch.zeros((len(names) + 2, b), dtype=torch.float32, device=dev)   # distinct rows: no aliasing
    reward, done = zeros[0], zeros[1]
    met = {k: zeros[2 + i] for i, k in enumerate(names)}
    if batch is None:
      qp = _system.QP(*[getattr(qp, k)[0] for k in ('pos', 'rot', 'vel', 'ang')])
      obs, reward, done, met = obs[0], reward[0], done[0], {k: v[0] for k, v in met.items()}
    return State(qp, obs, reward, done, met)

  def _place_targets(self, qp, gen):
    return qp

  def step(self, state: State, action, host_out=None) -> State:
    """`Env.step(state, action)` (ant.py:222, humanoid.py:246): one kernel launch.

    host_out: (state on the GPU) the dict of `host_io(batch)`: obs / reward / done / metrics are
    written straight into those page-locked host tensors by the launch, and `action` may be its
    page-locked 'action' tensor -- the gym-style loop (wrappers.py:228-260, to_torch.py:56-64) with
    the state resident on the device.  Synchronise before reading them."""
    return self._launch(state, action, host_out=host_out)

  def randomize_friction(self, friction):
    """Domain randomisation over collider friction, as `DomainRandomizationWrapper(env_fn,
    *friction_randomizer(env, friction))` (experimental/tracing/wrappers.py:27-88,
    randomizers.py:26-51): env e is simulated with material.friction = friction[e] on EVERY collider
    (a pair's friction is friction[e]^2, colliders.py:277,315,392).  friction: float32 CUDA tensor [B] matching
    the batch of the states stepped afterwards, or None to return to the config's friction.  Applies to
    `step`, `reset` (the reset obs reads System.info's contact impulse) and the wrappers around this env."""
    self.friction = friction
    return self

  def randomize_joints(self, joint_offset):
    """Domain randomisation over joint sockets, as `DomainRandomizationWrapper(env_fn, *joint_randomizer(env,
    offsets, keys))` (experimental/tracing/randomizers.py:54-97): env e is stepped with `Joint.parent_offset` =
    joint_offset[e, j] for joint j -- `self.sys.joint_offsets()` (the config's values) plus your per-env deltas on
    the joints you select.  float32 CUDA tensor [B,J,3], or None to return to the config.  Applies to `step`, to
    `reset` (the pose is built from the randomised sockets, as the reference's vmapped `default_qp` does:
    bx_default_qp_offsets) and to the wrappers around this env."""
    self.joint_offset = joint_offset
    return self

  def host_io(self, batch: int):
    """Page-locked host tensors for `step(..., host_out=)`: 'action' [B,A] in, 'obs' [B,O],
    'reward' [B], 'done' [B], 'metrics' [B,M] out."""
    import torch
    mk = lambda *s: torch.zeros(s, dtype=torch.float32).pin_memory()  # noqa: E731
    return {'action': mk(batch, self.action_size), 'obs': mk(batch, self.observation_size), 'reward': mk(batch),
            'done': mk(batch), 'metrics': mk(batch, self.sys.num_metrics)}

  def _launch(self, state: State, action, rand=None, auto_reset=None, host_out=None) -> State:
    torch_path = _system._is_torch(state.qp.pos)
    qp, info, eo = self.sys.step(state.qp, action, env=True, env_rand=rand, auto_reset=auto_reset,
                                 info=self.keep_info or not torch_path, friction=self.friction, joint_offset=self.joint_offset,
                                 env_out=None if host_out is None else {k: host_out[k] for k in ('obs', 'reward', 'done', 'metrics')})
    names = self.env.metric_names()
    metrics = dict(state.metrics)
    metrics.update({k: eo['metrics'][..., i] for i, k in enumerate(names)})
    extra = {k: eo[k] for k in ('steps', 'truncation') if k in eo}
    if info is not None:
      extra['physics'] = info
    return State(qp, eo['obs'], eo['reward'], eo['done'], metrics, dict(state.info, **extra))


class Ant(Env):
  """brax/envs/ant.py:26-281 (constructor arguments and defaults :173-196)."""

  def __init__(self, ctrl_cost_weight=0.5, use_contact_forces=False, contact_cost_weight=5e-4,
               healthy_reward=1.0, terminate_when_unhealthy=True, healthy_z_range=(0.2, 1.0),
               reset_noise_scale=0.1, exclude_current_positions_from_observation=True,
               legacy_spring=False, device: int = 0):
    self._reset_noise_scale = reset_noise_scale
    super().__init__('ant_spring' if legacy_spring else 'ant', EnvDesc(
        'ant', 1.0, ctrl_cost_weight, contact_cost_weight, healthy_reward, terminate_when_unhealthy,
        tuple(healthy_z_range), use_contact_forces, exclude_current_positions_from_observation), device)


class Humanoid(Env):
  """brax/envs/humanoid.py:26-340 (constructor arguments and defaults :200-221)."""

  def __init__(self, forward_reward_weight=1.25, ctrl_cost_weight=0.1, healthy_reward=5.0,
               terminate_when_unhealthy=True, healthy_z_range=(0.8, 2.1), reset_noise_scale=1e-2,
               exclude_current_positions_from_observation=True, legacy_spring=False, device: int = 0):
    self._reset_noise_scale = reset_noise_scale
    super().__init__('humanoid_spring' if legacy_spring else 'humanoid', EnvDesc(
        'humanoid', forward_reward_weight, ctrl_cost_weight, 0.0, healthy_reward, terminate_when_unhealthy,
        tuple(healthy_z_range), False, exclude_current_positions_from_observation), device)


class _TargetEnv(Env):
  """Common part of fetch / grasp: targets that were hit are teleported to a new random place
  (fetch.py:88-93 + :124-134, grasp.py:131-136 + :188-198) INSIDE the launch, from three
  uniform numbers per env handed to it (`bx_step_io.env_rand`).  The reference draws them from
  the JAX PRNG carried in `state.info['rng']`; here a numpy / torch generator (seeded at
  construction) does -- same distribution, not the same stream; pass `rand=` to `step` to
  supply your own (e.g. the reference's draws)."""

  def __init__(self, system_name, env, device=0, seed=0):
    super().__init__(system_name, env, device)
    self._np_rng = np.random.default_rng(seed)
    self._seed = seed
    self._torch_gen = None

  def _uniform(self, like, lead):
    if _system._is_torch(like):
      import torch
      if self._torch_gen is None or self._torch_gen.device != like.device:
        self._torch_gen = torch.Generator(device=like.device)
        self._torch_gen.manual_seed(self._seed)
      return torch.rand(tuple(lead) + (3,), generator=self._torch_gen, device=like.device, dtype=torch.float32)
    return self._np_rng.uniform(size=tuple(lead) + (3,)).astype(np.float32)

  def step(self, state: State, action, rand=None) -> State:
    return self._launch(state, action, rand)

  def _launch(self, state: State, action, rand=None, auto_reset=None) -> State:
    if rand is None:
      rand = self._uniform(state.qp.pos, state.qp.pos.shape[:-2])
    return super()._launch(state, action, rand, auto_reset)


class Fetch(_TargetEnv):
  """brax/envs/fetch.py:25-134."""

  def __init__(self, legacy_spring=False, device: int = 0, seed: int = 0):
    super().__init__('fetch_spring' if legacy_spring else 'fetch', fetch_desc, device, seed)

  def _place_targets(self, qp, gen):
    # reset's `_random_target` (fetch.py:38-40, :124-134)
    import math
    import torch
    b = qp.pos.shape[0]
    u = torch.rand(b, 2, generator=gen, device=qp.pos.device, dtype=torch.float32)
    dist = self.env.target_radius + self.env.target_distance * u[:, 0]
    ang = (math.pi * 2.) * u[:, 1]
    qp.pos[:, self.env.target_body] = torch.stack([dist * torch.cos(ang), dist * torch.sin(ang), torch.ones_like(dist)], -1)
    return qp


class Grasp(_TargetEnv):
  """brax/envs/grasp.py:25-198 (pbd; action_size = actuator dofs + 3 palm-target slots :140-142)."""

  def __init__(self, legacy_spring=False, device: int = 0, seed: int = 0):
    if legacy_spring:
      raise NotImplementedError('grasp env epilogue: pbd only (see grasp_desc)')
    super().__init__('grasp', grasp_desc, device, seed)


class EpisodeAutoReset:
  """`AutoResetWrapper(EpisodeWrapper(env, episode_length, 1))` (brax/envs/wrappers.py:95-148,
  stacked as in brax/envs/__init__.py:84-89) with both wrappers' `step` evaluated inside the env's
  single kernel launch (`bx_step_io.first_* / done_in / steps_in / episode_length`): episode
  counter, truncation, done at the episode limit, and the reload of (qp, obs) from the first
  state where done.  torch CUDA states only.  episode_length None / 0: AutoResetWrapper alone."""

  def __init__(self, env: Env, episode_length: int = 1000, action_repeat: int = 1, resample: bool = False):
    """action_repeat: EpisodeWrapper's (wrappers.py:106-111): every step runs that many env steps on
    the same action inside the launch and returns the sum of their rewards.
    resample: draw a NEW noisy first state (and its obs) before every step / rollout instead
    of reusing the one `reset` stored -- per-episode reset noise (the reference
    restarts every episode of an env from the same first state, wrappers.py:146-147).  Costs the
    two reset launches (bx_default_qp + observe) per call."""
    self.env = env
    self.sys = env.sys
    self.resample = bool(resample)
    self.action_repeat = max(1, int(action_repeat or 1))
    self._draws = 0
    self.episode_length = int(episode_length or 0)
    self.observation_size = env.observation_size
    self.action_size = env.action_size

  def reset(self, *args, first_obs=None, **kw) -> State:
    """As `env.reset`, plus info['steps'], info['truncation'] (wrappers.py:98-103) and
    info['first_qp'], info['first_obs'] (wrappers.py:126-129): the reset state and its obs
    (`first_obs=` overrides the latter)."""
    st = self.env.reset(*args, **kw)
    if not _system._is_torch(st.qp.pos):
      raise ValueError('EpisodeAutoReset needs torch CUDA states: reset(batch, device=...)')
    import torch
    if first_obs is not None:
      st = st.replace(obs=torch.as_tensor(first_obs, dtype=torch.float32, device=st.obs.device).expand_as(st.obs).contiguous())
    st.info.update(steps=torch.zeros_like(st.done), truncation=torch.zeros_like(st.done),
                   first_qp=st.qp, first_obs=st.obs)
    return st

  def _auto_reset(self, state: State):
    first_qp, first_obs = state.info['first_qp'], state.info['first_obs']
    if self.resample:
      self._draws += 1
      fresh = self.env.reset(int(state.qp.pos.shape[0]), device=state.qp.pos.device, seed=1_000_003 * self._draws)
      first_qp, first_obs = fresh.qp, fresh.obs
    return _system.AutoReset(first_qp, first_obs, state.done, state.info.get('steps'), self.episode_length,
                             self.action_repeat)

  def step(self, state: State, action, rand=None) -> State:
    if rand is None and self.action_repeat > 1 and isinstance(self.env, _TargetEnv):
      rand = self.env._uniform(state.qp.pos, (self.action_repeat,) + tuple(state.qp.pos.shape[:-2]))
    return self.env._launch(state, action, rand, self._auto_reset(state))

  def rollout(self, state: State, actions, rand=None):
    """T wrapped steps in ONE launch (the scan over `step` of a training / eval unroll).  Returns
    the final state and a dict of the per-step obs / reward / done / truncation / metrics."""
    ar = self._auto_reset(state)
    if rand is None and isinstance(self.env, _TargetEnv):
      rand = self.env._uniform(state.qp.pos, (int(actions.shape[0]) * self.action_repeat,) + tuple(state.qp.pos.shape[:-2]))
    qp, info, eo = self.sys.rollout(state.qp, actions, env=True, env_rand=rand, auto_reset=ar, info=False,
                                    friction=self.env.friction, joint_offset=self.env.joint_offset)
    names = self.env.env.metric_names()
    metrics = {k: eo['metrics'][-1, ..., i] for i, k in enumerate(names)}
    last = State(qp, eo['obs'][-1], eo['reward'][-1], eo['done'][-1], metrics,
                 dict(state.info, steps=eo['steps'], truncation=eo['truncation'][-1]))
    return last, eo


@dataclasses.dataclass
class EvalMetrics:
  """brax/envs/wrappers.py:150-165."""
  episode_metrics: Dict[str, Any]
  active_episodes: Any
  episode_steps: Any


class EvalWrapper:
  """`EvalWrapper` (brax/envs/wrappers.py:168-200) around an `EpisodeAutoReset`: per-env sums of
  the metrics (and the reward) over the first episode, its length and the still-active flag, kept
  in `state.info['eval_metrics']`.  `step` is the wrapped env's single launch plus three small
  elementwise updates; `rollout` folds the recurrence over the T steps of ONE launch."""

  def __init__(self, env: 'EpisodeAutoReset'):
    self.env = env
    self.sys = env.sys
    self.observation_size, self.action_size = env.observation_size, env.action_size

  def reset(self, *args, **kw) -> State:
    import torch
    st = self.env.reset(*args, **kw)
    st.metrics['reward'] = st.reward                                   # wrappers.py:173
    st.info['eval_metrics'] = EvalMetrics({k: torch.zeros_like(v) for k, v in st.metrics.items()},
                                          torch.ones_like(st.reward), torch.zeros_like(st.reward))
    return st

  def step(self, state: State, action, rand=None) -> State:
    import torch
    em = state.info['eval_metrics']
    if not isinstance(em, EvalMetrics):
      raise ValueError('Incorrect type for state_metrics: %s' % type(em))
    n = self.env.step(state, action, rand)
    n.metrics['reward'] = n.reward                                     # wrappers.py:185
    act = em.active_episodes
    n.info['eval_metrics'] = EvalMetrics(
        {k: em.episode_metrics[k] + n.metrics[k] * act for k in n.metrics},          # :189-191
        act * (1 - n.done),                                                           # :192
        torch.where(act > 0, n.info['steps'], em.episode_steps))                      # :187-188
    return n

  def rollout(self, state: State, actions, rand=None):
    """T wrapped steps in one launch; the eval bookkeeping of all T steps from the [T,B] outputs."""
    import torch
    em = state.info['eval_metrics']
    s0 = torch.where(state.done > 0, torch.zeros_like(state.done), state.info['steps'])
    last, eo = self.env.rollout(state, actions, rand)
    alive = torch.cumprod(1 - eo['done'], 0)                           # active after step t, given active before
    before = torch.cat([torch.ones_like(alive[:1]), alive[:-1]], 0) * em.active_episodes   # active before step t
    names = self.env.env.env.metric_names()
    sums = {k: (eo['metrics'][..., i] * before).sum(0) for i, k in enumerate(names)}
    sums['reward'] = (eo['reward'] * before).sum(0)
    last.metrics['reward'] = last.reward
    # while an env is active nothing has reset its counter: steps after wrapped step t =
    # s0 + action_repeat * (t + 1) (EpisodeWrapper advances `steps` by action_repeat, wrappers.py:112)
    last.info['eval_metrics'] = EvalMetrics(
        {k: em.episode_metrics[k] + sums[k] for k in sums},
        em.active_episodes * alive[-1],
        torch.where(em.active_episodes > 0, s0 + float(self.env.action_repeat or 1) * before.sum(0),
                    em.episode_steps))
    return last, eo


_ENVS = {'ant': Ant, 'humanoid': Humanoid, 'fetch': Fetch, 'grasp': Grasp}


def create(env_name: str, episode_length: int = None, action_repeat: int = 1, auto_reset: bool = False,
           eval_metrics: bool = False, **kwargs):
  """`envs.create(name, episode_length, auto_reset=...)` (brax/envs/__init__.py:74-90) for the
  mirrored envs.  'ant' is registered with use_contact_forces=True as in this fork
  (brax/envs/__init__.py:45).  Unlike the reference the wrappers are opt-in (defaults: the bare
  env); with auto_reset=True the result is `EpisodeAutoReset(env, episode_length)`."""
  if env_name not in _ENVS:
    raise KeyError('env %r has no fused epilogue (available: %s)' % (env_name, sorted(_ENVS)))
  if env_name == 'ant':
    kwargs.setdefault('use_contact_forces', True)
  env = _ENVS[env_name](**kwargs)
  if auto_reset:
    env = EpisodeAutoReset(env, episode_length, action_repeat)
    return EvalWrapper(env) if eval_metrics else env      # brax/envs/__init__.py:88-89
  if eval_metrics:
    raise ValueError('eval_metrics needs auto_reset=True (EvalWrapper reads info[\'steps\'])')
  if episode_length or (action_repeat or 1) > 1:
    raise ValueError('episode_length / action_repeat without auto_reset are not fused: wrap the env yourself')
  return env
