"""bx_default_qp (System.default_qp(joint_angle, joint_velocity), system.py:112-242, as one kernel
launch) against the reference's own outputs (tests/golden/reset.npz) and the numpy oracle."""
import os

import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(util.GOLDEN, 'reset.npz'))
NAMES = sorted({k.split('/')[0] for k in G.files})


def _torch():
  import torch
  if not torch.cuda.is_available():
    pytest.skip('no CUDA device')
  return torch


@pytest.mark.parametrize('name', NAMES)
def test_default_qp_matches_reference(name):
  torch = _torch()
  from brax_b200 import system as bsys
  s = bsys.load_system(name)
  assert s.num_joint_dof == G[name + '/angle'].shape[1]
  ja = torch.from_numpy(G[name + '/angle'].astype(np.float32)).cuda()
  jv = torch.from_numpy(G[name + '/velocity'].astype(np.float32)).cuda()
  qp = s.default_qp(joint_angle=ja, joint_velocity=jv)
  torch.cuda.synchronize()
  for k in util.QP_KEYS:
    np.testing.assert_allclose(getattr(qp, k).cpu().numpy(), G[name + '/' + k], rtol=1e-5, atol=3e-6, err_msg=k)
  # default_qp() itself: the recorded default state
  q0 = s.default_qp(joint_angle=torch.from_numpy(s.default_angle()).cuda())
  assert q0.pos.shape == (s.num_bodies, 3)
  for k in util.QP_KEYS:
    np.testing.assert_allclose(getattr(q0, k).cpu().numpy(), np.asarray(s.desc.default_qp[k]), rtol=1e-5, atol=3e-6)


@pytest.mark.parametrize('name,b', [('ant', 4133), ('humanoid', 2048), ('grasp', 1000), ('fetch_spring', 777)])
def test_default_qp_matches_oracle_on_large_batches(name, b):
  """Ragged batches, noise as in the envs' reset and wider; float32 oracle on the same inputs."""
  torch = _torch()
  from oracle import reset_oracle
  from brax_b200 import system as bsys
  s = bsys.load_system(name)
  rng = np.random.default_rng(5)
  ja = (s.default_angle() + rng.uniform(-.6, .6, (b, s.num_joint_dof))).astype(np.float32)
  jv = rng.uniform(-1, 1, (b, s.num_joint_dof)).astype(np.float32)
  qp = s.default_qp(joint_angle=torch.from_numpy(ja).cuda(), joint_velocity=torch.from_numpy(jv).cuda())
  want = reset_oracle.default_qp(s.desc, ja, jv, np.float32)
  w64 = reset_oracle.default_qp(s.desc, ja, jv, np.float64)
  for k in util.QP_KEYS:
    got = getattr(qp, k).cpu().numpy()
    np.testing.assert_allclose(got, want[k], rtol=1e-5, atol=3e-6, err_msg=k)
    np.testing.assert_allclose(got, w64[k], rtol=1e-5, atol=5e-6, err_msg=k)
  # the lowest probe of every lowered tree touches the plane
  z = qp.pos[..., 2].cpu().numpy()
  assert z.min() > -1e-5
  # velocities alone / angles alone
  q1 = s.default_qp(joint_velocity=torch.from_numpy(jv).cuda())
  w1 = reset_oracle.default_qp(s.desc, np.tile(s.default_angle(), (b, 1)), jv, np.float32)
  np.testing.assert_allclose(q1.ang.cpu().numpy(), w1['ang'], rtol=1e-5, atol=3e-6)


def test_default_qp_argument_errors():
  torch = _torch()
  from brax_b200 import system as bsys
  s = bsys.load_system('ant')
  with pytest.raises(RuntimeError):
    s.default_qp(joint_angle=torch.zeros(4, s.num_joint_dof))          # CPU tensor: no fallback
  with pytest.raises(RuntimeError):
    s.default_qp(joint_angle=torch.zeros(4, s.num_joint_dof + 1).cuda())


def _touching(desc, ins):
  """Envs with a contact row whose penetration is within rounding of zero (a collider resting exactly ON
  the plane, as after default_qp's lowering): `penetration > 0` -- and with it the whole contact impulse
  of System.info(qp), colliders.py:131-133 -- is a coin flip per float32 evaluation, in the reference too."""
  from oracle.oracle import Oracle
  pen = Oracle(desc, np.float64).info_contact(ins['pos'], ins['rot'], ins['vel'], ins['ang'])['contact_penetration']
  return (np.abs(pen) < 1e-5).any(axis=1) if pen.size else np.zeros(ins['pos'].shape[0], bool)


def _contact_block(env, desc):
  """Indices of the obs entries that read info.contact: ant's clipped contact forces (ant.py:271-279),
  fetch / grasp's contact flags (fetch.py:116-117, grasp.py:176-178); humanoid's obs reads none."""
  n, o = desc.num_bodies, env.obs_size(desc)
  if env.kind == 'ant':
    return np.arange(o - 6 * n, o) if env.use_contact_forces else np.arange(0)
  if env.kind in ('fetch', 'grasp'):
    return np.arange(o - n, o)
  return np.arange(0)


def _obs_close(obs, want, tol, touching, block):
  """Equal to the obs built on sys.info(qp), except that a `touching` env's contact entries are exempt
  (each of its resting contacts may be decided either way by either evaluation)."""
  ok = np.abs(obs - want) <= tol
  exempt = np.zeros_like(ok)
  exempt[np.ix_(np.nonzero(touching)[0], block)] = True
  return ok | exempt


@pytest.mark.parametrize('variant,b,steps', [('ant', 1024, 25), ('ant_spring', 1024, 25), ('humanoid', 1024, 25),
                                             ('fetch', 1024, 15), ('grasp', 1024, 15)])
def test_observe_info_matches_oracle_on_moving_states(variant, b, steps):
  """System.info(qp) (system.py:327-340 / :378-390) as the observe launch evaluates it: info.contact and
  the contact rows on states taken mid-simulation (bodies pressing into the ground with velocity, so the
  contacts are decided well away from rounding) against the oracle's `info_contact`, pinned to the
  reference's sys.info(qp) by tests/test_env_oracle.py."""
  torch = _torch()
  from oracle.oracle import Oracle
  from tests.test_gpu_env import _env
  from tests.test_gpu_parity import _harvest
  sysname, env = _env(variant)
  qp, _ = _harvest(env.sys, b, steps, 17)
  # a step leaves its contacts resolved (normal velocity >= 0: no impulse): push the free bodies down
  free = torch.tensor([float(min(x['pos_mask'])) for x in env.sys.desc.bodies], device='cuda')
  vel = qp.vel.clone()
  vel[..., 2] -= 0.3 * free
  qp = type(qp)(qp.pos, qp.rot, vel, qp.ang)
  _, info, eo = env.sys._step_torch(qp, torch.zeros(b, env.action_size).cuda(), 1, False, True, None, None, True,
                                    observe=True)
  ins = {k: getattr(qp, k).cpu().numpy() for k in util.QP_KEYS}
  o32 = Oracle(env.sys.desc, np.float32).info_contact(ins['pos'], ins['rot'], ins['vel'], ins['ang'])
  o64 = Oracle(env.sys.desc, np.float64).info_contact(ins['pos'], ins['rot'], ins['vel'], ins['ang'])
  got = {'contact_vel': info.contact.vel, 'contact_ang': info.contact.ang, 'contact_pos': info.contact_pos,
         'contact_normal': info.contact_normal, 'contact_penetration': info.contact_penetration}
  live = np.abs(o32['contact_vel']).reshape(b, -1).max(axis=1) > 0
  assert live.mean() > 0.05, live.mean()                 # the impulse is exercised
  edge = (np.abs(o64['contact_penetration']) < 1e-5).any(axis=1)   # envs with a contact at rounding of 0
  bad = np.zeros(b, bool)
  for k, v in got.items():
    g = v.cpu().numpy()
    assert np.isfinite(g).all(), k
    tol = 2e-6 + 1e-5 * np.abs(o32[k]) + 4 * np.abs(o32[k] - o64[k])
    if k in ('contact_vel', 'contact_ang'):
      tol = tol + 1e-5 * np.abs(o32[k]).reshape(b, -1).max(axis=1).reshape((b,) + (1,) * (g.ndim - 1))
    bad |= (np.abs(g - o32[k]) > tol).reshape(b, -1).any(axis=1)
  assert (bad & ~edge).sum() <= int(0.005 * b), int((bad & ~edge).sum())
  flags = (got['contact_penetration'].cpu().numpy() > 0) == (o32['contact_penetration'] > 0)
  assert (flags | (np.abs(o32['contact_penetration']) < 1e-6)).all()


@pytest.mark.parametrize('variant', ['ant', 'ant_spring', 'humanoid', 'humanoid_spring', 'fetch', 'grasp'])
def test_observe_matches_reference_reset_obs(variant):
  """System.observe (the env epilogue without substeps, on `sys.info(qp)` of the state) on the golden
  reset states against the obs of the reference's own `_get_obs(qp, sys.info(qp))` (reset_obs.npz)
  on EVERY draw -- including the humanoid / fetch draws whose info.contact is non-zero -- and against
  the float32 oracle."""
  torch = _torch()
  from oracle import env_oracle
  from oracle.oracle import Oracle
  from tests.test_gpu_env import _env
  from brax_b200.system import QP
  go = np.load(os.path.join(util.GOLDEN, 'reset_obs.npz'))
  sysname, env = _env(variant)
  ins = {k: G[sysname + '/' + k].astype(np.float32) for k in util.QP_KEYS}
  qp = QP(*[torch.from_numpy(ins[k]).cuda() for k in util.QP_KEYS])
  obs = env.sys.observe(qp).cpu().numpy()
  assert obs.shape == (ins['pos'].shape[0], env.observation_size)
  want = env_oracle.reset_obs(env.sys.desc, env.env, Oracle(env.sys.desc, np.float32), ins, np.float32)
  # 1e-5 relative + the float32 conditioning of the obs terms that cancel (humanoid's com_ang)
  w64 = env_oracle.reset_obs(env.sys.desc, env.env, Oracle(env.sys.desc, np.float64), ins, np.float64)
  tol = 5e-6 + 1e-5 * np.abs(want) + 4 * np.abs(want - w64)
  touching = _touching(env.sys.desc, ins)
  block = _contact_block(env.env, env.sys.desc)
  assert _obs_close(obs, want, tol, touching, block).all()
  # ... and the reference's own obs (float64), on every draw.  After default_qp the lowest collider of
  # every reset state rests exactly ON the plane, so whether that contact counts is rounding-level in
  # the reference itself; the contact entries of such states are compared on moving states instead
  # (test_observe_info_matches_oracle_on_moving_states).
  ref = go[variant + '/obs']
  assert _obs_close(obs, ref, 2e-5 + 1e-5 * np.abs(ref), touching, block).all()
  # the state passes through untouched; unbatched call == row 2
  q2, _, eo = env.sys._step_torch(qp, torch.zeros(obs.shape[0], env.action_size).cuda(), 1, False, True, None, None,
                                  False, observe=True)
  for k in util.QP_KEYS:
    assert torch.equal(getattr(q2, k), getattr(qp, k)), k
  one = env.sys.observe(QP(*[getattr(qp, k)[2] for k in util.QP_KEYS]))
  assert torch.equal(one, torch.from_numpy(obs[2]).cuda())


@pytest.mark.parametrize('variant', ['ant', 'humanoid', 'fetch', 'grasp'])
def test_env_reset_on_the_gpu(variant):
  """Env.reset(batch, device='cuda'): the reference's recipe -- noisy joint state -> default_qp
  (forward kinematics) -> obs -- against the oracles on the joint state that was drawn."""
  torch = _torch()
  from oracle import env_oracle, reset_oracle
  from oracle.oracle import Oracle
  from tests.test_gpu_env import _env
  sysname, env = _env(variant)
  b = 333
  st = env.reset(b, device='cuda', seed=3)
  st2 = env.reset(b, device='cuda', seed=3)
  assert torch.equal(st.qp.pos, st2.qp.pos) and torch.equal(st.obs, st2.obs)
  assert st.obs.shape == (b, env.observation_size) and float(st.reward.abs().max()) == 0 and float(st.done.max()) == 0
  assert set(st.metrics) == set(env.env.metric_names())
  desc = env.sys.desc
  scale = getattr(env, '_reset_noise_scale', 0.0)
  # redraw the joint state with the same generator to check the kernel chain on it
  gen = torch.Generator(device='cuda'); gen.manual_seed(3)
  nd = env.sys.num_joint_dof
  ja = torch.from_numpy(env.sys.default_angle()).cuda() + (torch.rand(b, nd, generator=gen, device='cuda') * 2 - 1) * scale
  jv = (torch.rand(b, nd, generator=gen, device='cuda') * 2 - 1) * scale
  want = reset_oracle.default_qp(desc, ja.cpu().numpy(), jv.cpu().numpy(), np.float32)
  pos = st.qp.pos.cpu().numpy()
  if env.env.kind == 'fetch':     # the random target: on the circle of fetch.py:124-134, z = 1
    t = pos[:, env.env.target_body]
    r = np.hypot(t[:, 0], t[:, 1])
    assert (r >= env.env.target_radius - 1e-4).all() and (r <= env.env.target_radius + env.env.target_distance + 1e-4).all()
    assert np.allclose(t[:, 2], 1.0) and np.std(np.arctan2(t[:, 1], t[:, 0])) > 1.0
    want['pos'][:, env.env.target_body] = t
  for k in util.QP_KEYS:
    np.testing.assert_allclose(getattr(st.qp, k).cpu().numpy(), want[k], rtol=1e-5, atol=3e-6, err_msg=k)
  if scale:
    assert np.std(pos[:, 1, 2]) > 1e-4                      # per-env noise reached the poses
  ins = {k: getattr(st.qp, k).cpu().numpy() for k in util.QP_KEYS}
  wobs = env_oracle.reset_obs(desc, env.env, Oracle(desc, np.float32), ins, np.float32)
  w64 = env_oracle.reset_obs(desc, env.env, Oracle(desc, np.float64), ins, np.float64)
  tol = 5e-6 + 1e-5 * np.abs(wobs) + 4 * np.abs(wobs - w64)
  assert _obs_close(st.obs.cpu().numpy(), wobs, tol, _touching(desc, ins), _contact_block(env.env, desc)).all()
  # caller-supplied joint state; unbatched reset
  st3 = env.reset(b, device='cuda', joint_angle=ja, joint_velocity=jv)
  np.testing.assert_allclose(st3.qp.rot.cpu().numpy(), want['rot'], rtol=1e-5, atol=3e-6)
  one = env.reset(device='cuda', seed=1)
  assert one.qp.pos.shape == (desc.num_bodies, 3) and one.obs.shape == (env.observation_size,)
  # and the state steps: a wrapped rollout from it stays finite
  from brax_b200 import envs
  w = envs.EpisodeAutoReset(env, 5)
  s0 = w.reset(b, device='cuda', seed=2)
  assert torch.equal(s0.info['first_obs'], s0.obs)
  acts = torch.rand(12, b, env.action_size, device='cuda') * 2 - 1
  last, eo = w.rollout(s0, acts)
  assert bool(torch.isfinite(eo['obs']).all()) and float(eo['truncation'].sum()) > 0
  done = eo['done'][4] > 0                 # episode_length 5: envs without an earlier done end at step 5
  early = eo['done'][:4].sum(0) > 0
  assert bool((done | early).all()) and float(done.float().mean()) > 0.5
  assert torch.equal(eo['obs'][4][done], s0.obs[done])


def test_resampled_auto_reset_draws_new_first_states():
  """EpisodeAutoReset(resample=True): every reset inside the launch starts from a freshly drawn
  noisy state (two reset launches per call), not from the state `reset` stored."""
  torch = _torch()
  from brax_b200 import envs
  from tests.test_gpu_env import _env
  _, env = _env('ant')
  b = 256
  w = envs.EpisodeAutoReset(env, 2, resample=True)
  st = w.reset(b, device='cuda', seed=0)
  act = torch.zeros(b, env.action_size, device='cuda')
  st = w.step(st, act)
  st = w.step(st, act)                      # steps == 2: every env is reset ...
  assert bool((st.done > 0).all())
  assert not torch.equal(st.qp.pos, st.info['first_qp'].pos)   # ... to a state that is not the stored one
  z0 = st.qp.pos[:, 0, 2]
  assert float((z0 - st.info['first_qp'].pos[:, 0, 2]).abs().max()) < 0.2   # but a reset pose all the same
  st = w.step(st, act)
  st2 = w.step(st, act)
  assert bool((st2.done > 0).all()) and not torch.equal(st2.qp.pos, st.qp.pos)
