"""Per-env collider friction (domain randomisation; `bx_step_io.friction`, `System.step(friction=)`).

Reference: `DomainRandomizationWrapper` + `friction_randomizer` (/root/reference/brax/experimental/tracing/
wrappers.py:27-88, randomizers.py:26-51): env e steps a System whose colliders all have material.friction =
friction[e].  tests/golden/friction/*.npz holds such steps produced by the unmodified reference through its own
TracedConfig hook, one System per value (oracle/gen_golden_friction.py).

CPU: the float64 oracle with the substituted pair frictions reproduces them.  GPU: ONE launch over a batch whose
envs carry different friction values matches the float32 oracle run per value."""
import os

import numpy as np
import pytest

from tests import util

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'friction')
SYSTEMS = ('ant', 'ant_spring', 'grasp')


def _golden(name):
  g = np.load(os.path.join(GOLD, name + '.npz'))
  f, t = g['pos'].shape[:2]
  flat = {k: g[k].reshape((f * t,) + g[k].shape[2:]) for k in g.files if k not in ('friction', 'pair_friction')}
  friction = np.repeat(g['friction'], t).astype(np.float32)
  a = util.load_desc(name).act_size      # (grasp: the descriptor's act vector carries the env's 3 palm-target slots too)
  if flat['act'].shape[1] < a:
    flat['act'] = np.concatenate([flat['act'], np.zeros((f * t, a - flat['act'].shape[1]))], axis=1)
  return g, flat, friction


@pytest.mark.parametrize('name', SYSTEMS)
def test_oracle_f64_reproduces_reference_steps_under_substituted_friction(name):
  from oracle.oracle import FrictionOracle
  g, flat, friction = _golden(name)
  # the extractor read friction_a = friction_b = the substituted value off every built System
  assert g['pair_friction'].shape[1] == len(util.load_desc(name).pairs)
  for i, f in enumerate(g['friction']):
    assert np.all(g['pair_friction'][i] == f)
  t = g['pos'].shape[1]
  o = FrictionOracle(util.load_desc(name), np.repeat(g['friction'], t), np.float64)
  out = o.step(*[flat[k] for k in util.QP_KEYS], flat['act'])
  for k in util.QP_KEYS + util.INFO_KEYS:
    if out[k].size:
      np.testing.assert_allclose(out[k], flat['out_' + k], rtol=0, atol=1e-9 * max(1.0, np.abs(flat['out_' + k]).max()), err_msg=k)
  # and friction matters in these states: the slipperiest and the stickiest lane end up apart
  assert np.abs(g['out_pos'][0, -1] - g['out_pos'][-1, -1]).max() > 1e-2


def test_friction_oracle_equals_plain_oracle_at_the_config_value():
  from oracle.oracle import FrictionOracle, Oracle
  d = util.load_desc('ant')
  f0 = d.pairs[0]['friction_a']
  assert all(p['friction_a'] == f0 and p['friction_b'] == f0 for p in d.pairs)
  g = util.load_golden('ant')
  ins = [g[k].astype(np.float32) for k in util.QP_KEYS]
  act = g['act'].astype(np.float32)
  a = Oracle(d, np.float32).step(*ins, act)
  b = FrictionOracle(d, np.full(ins[0].shape[0], f0, np.float32), np.float32).step(*ins, act)
  for k in a:
    assert np.array_equal(a[k], b[k]), k


# ------------------------------------------------------------------ GPU
def _torch():
  import torch
  if not torch.cuda.is_available():
    pytest.skip('no CUDA device')
  return torch


@pytest.mark.gpu
@pytest.mark.parametrize('name', SYSTEMS)
def test_gpu_step_with_per_env_friction_on_reference_states(name):
  torch = _torch()
  from oracle.oracle import FrictionOracle
  from brax_b200.system import QP
  from tests.test_gpu_parity import _flat, _system, check_step
  g, flat, friction = _golden(name)
  s = _system(name)
  ins = [flat[k].astype(np.float32) for k in util.QP_KEYS]
  act = flat['act'].astype(np.float32)
  qp = QP(*[torch.from_numpy(x).cuda() for x in ins])
  got = _flat(*s.step(qp, torch.from_numpy(act).cuda(), friction=torch.from_numpy(friction).cuda()))
  d = util.load_desc(name)
  check_step(name, ins, act, got, oracles=(FrictionOracle(d, friction, np.float32), FrictionOracle(d, friction, np.float64)))
  # and it is the friction that moved them: the same launch without it differs on the non-config lanes
  plain = _flat(*s.step(qp, torch.from_numpy(act).cuda()))
  lanes = friction != np.float32(d.pairs[0]['friction_a'])
  assert np.abs(plain['vel'][lanes] - got['vel'][lanes]).max() > 1e-3
  assert np.array_equal(plain['pos'][~lanes], got['pos'][~lanes])


@pytest.mark.gpu
@pytest.mark.parametrize('name,b', [('ant', 1000), ('grasp', 333), ('humanoid', 512), ('fetch', 300), ('ant_spring', 512)])
def test_gpu_per_env_friction_on_harvested_states(name, b):
  """A contact-rich batch, 8 friction values spread over the envs (ragged last tile), one launch."""
  torch = _torch()
  from oracle.oracle import FrictionOracle
  from tests.test_gpu_parity import _flat, _harvest, _system, check_step
  s = _system(name)
  qp, rng = _harvest(s, b, 6 if name == 'half_cheetah' else 20, 17)
  friction = rng.choice(np.float32([0.1, 0.3, 0.5, 0.75, 1.0, 1.25, 1.5, 2.0]), b).astype(np.float32)
  act = rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32)
  got = _flat(*s.step(qp, torch.from_numpy(act).cuda(), friction=torch.from_numpy(friction).cuda()))
  ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
  d = util.load_desc(name)
  check_step(name, ins, act, got, oracles=(FrictionOracle(d, friction, np.float32), FrictionOracle(d, friction, np.float64)))


@pytest.mark.gpu
def test_gpu_friction_at_the_config_value_is_bitwise_the_plain_step_and_rollouts_chain():
  torch = _torch()
  from tests.test_gpu_parity import _harvest, _system
  s = _system('ant')
  d = util.load_desc('ant')
  b, t_n = 700, 5
  qp, rng = _harvest(s, b, 20, 19)
  acts = torch.from_numpy(rng.uniform(-1, 1, (t_n, b, s.act_size)).astype(np.float32)).cuda()
  same = torch.full((b,), float(d.pairs[0]['friction_a']), device='cuda')
  q0, i0 = s.step(qp, acts[0])
  q1, i1 = s.step(qp, acts[0], friction=same)
  for k in util.QP_KEYS:
    assert torch.equal(getattr(q0, k), getattr(q1, k)), k
  assert torch.equal(i0.contact.vel, i1.contact.vel)
  # T steps in one launch == T single launches, with friction
  fr = torch.from_numpy(rng.uniform(0.2, 1.8, b).astype(np.float32)).cuda()
  qr, _ = s.rollout(qp, acts, friction=fr)
  qc = qp
  for t in range(t_n):
    qc, _ = s.step(qc, acts[t], friction=fr)
  for k in util.QP_KEYS:
    assert torch.equal(getattr(qr, k), getattr(qc, k)), k
  # the next launch without friction is back on the tables
  q2, _ = s.step(qp, acts[0])
  assert torch.equal(q2.pos, q0.pos)
  with pytest.raises((ValueError, RuntimeError)):
    s.step(qp, acts[0], friction=fr[:10])


@pytest.mark.gpu
def test_gpu_env_randomize_friction():
  """`Env.randomize_friction` = DomainRandomizationWrapper over friction: step, reset obs and the wrapped env."""
  torch = _torch()
  from brax_b200 import envs
  env = envs.create('ant')
  b = 256
  st = env.reset(b, device='cuda', seed=3)
  act = torch.rand(b, env.action_size, device='cuda') * 2 - 1
  for _ in range(15):
    st = env.step(st, act)
  plain = env.step(st, act)
  fr = torch.linspace(0.1, 2.0, b, device='cuda')
  env.randomize_friction(fr)
  rnd = env.step(st, act)
  q, _, eo = env.sys.step(st.qp, act, env=True, friction=fr, info=False)
  assert torch.equal(rnd.obs, eo['obs']) and torch.equal(rnd.qp.pos, q.pos)
  assert float((rnd.obs - plain.obs).abs().max()) > 1e-3
  wrapped = envs.EpisodeAutoReset(env, 1000)
  ws = wrapped.reset(b, device='cuda', seed=3)
  ws = ws.replace(qp=st.qp)
  w1 = wrapped.step(ws, act)
  assert torch.equal(w1.qp.pos, rnd.qp.pos)
  env.randomize_friction(None)
  assert torch.equal(env.step(st, act).obs, plain.obs)


# ------------------------------------------------------------------ per-env joint sockets (randomizers.py:54-97)
JOINT_SYSTEMS = ('ant', 'ant_spring', 'humanoid_spring')


def _golden_joints(name):
  g = np.load(os.path.join(GOLD, name + '_joints.npz'))
  v, t = g['pos'].shape[:2]
  flat = {k: g[k].reshape((v * t,) + g[k].shape[2:]) for k in g.files
          if k not in ('joint_offset', 'randomized') and not k.startswith('reset_')}
  off = np.repeat(g['joint_offset'], t, axis=0)          # [V*T, J, 3]
  return g, flat, off


@pytest.mark.parametrize('name', JOINT_SYSTEMS)
def test_oracle_f64_reproduces_reference_steps_under_substituted_joint_offsets(name):
  from oracle.oracle import FrictionOracle
  g, flat, off = _golden_joints(name)
  d = util.load_desc(name)
  # variant 0 is the config itself; the others moved exactly the randomised joints
  np.testing.assert_allclose(g['joint_offset'][0], [j['off_p'] for j in d.joints], atol=1e-7)
  moved = np.abs(g['joint_offset'][1] - g['joint_offset'][0]).max(axis=1) > 1e-4
  assert sorted(np.nonzero(moved)[0]) == sorted(g['randomized'])
  o = FrictionOracle(d, None, np.float64, joint_offset=off)
  out = o.step(*[flat[k] for k in util.QP_KEYS], flat['act'])
  for k in util.QP_KEYS + util.INFO_KEYS:
    if out[k].size:
      np.testing.assert_allclose(out[k], flat['out_' + k], rtol=0, atol=1e-9 * max(1.0, np.abs(flat['out_' + k]).max()), err_msg=k)
  assert np.abs(g['out_pos'][0, -1] - g['out_pos'][1, -1]).max() > 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize('name', JOINT_SYSTEMS)
def test_gpu_step_with_per_env_joint_offsets_on_reference_states(name):
  torch = _torch()
  from oracle.oracle import FrictionOracle
  from brax_b200.system import QP
  from tests.test_gpu_parity import _flat, _system, check_step
  g, flat, off = _golden_joints(name)
  s = _system(name)
  assert np.allclose(s.joint_offsets(), g['joint_offset'][0], atol=1e-7)
  ins = [flat[k].astype(np.float32) for k in util.QP_KEYS]
  act = flat['act'].astype(np.float32)
  qp = QP(*[torch.from_numpy(x).cuda() for x in ins])
  off32 = off.astype(np.float32)
  got = _flat(*s.step(qp, torch.from_numpy(act).cuda(), joint_offset=torch.from_numpy(off32).cuda()))
  d = util.load_desc(name)
  check_step(name, ins, act, got, oracles=(FrictionOracle(d, None, np.float32, joint_offset=off32),
                                           FrictionOracle(d, None, np.float64, joint_offset=off32)))
  # the lanes with the config's own sockets are bitwise the plain step; the others are not
  plain = _flat(*s.step(qp, torch.from_numpy(act).cuda()))
  t = g['pos'].shape[1]
  assert np.array_equal(plain['pos'][:t], got['pos'][:t]) and np.array_equal(plain['vel'][:t], got['vel'][:t])
  assert np.abs(plain['pos'][t:] - got['pos'][t:]).max() > 1e-4


@pytest.mark.gpu
def test_gpu_joint_offsets_with_friction_policy_and_rollout():
  """Both per-env tables and the in-kernel policy share the launch's extra shared memory: T steps in one launch ==
  T single launches; harvested ragged batch against the oracle per env."""
  torch = _torch()
  from oracle.oracle import FrictionOracle
  from brax_b200 import envs
  from brax_b200.system import Policy
  from tests.test_gpu_parity import _flat, _harvest, _system, check_step
  s = _system('humanoid')
  d = util.load_desc('humanoid')
  b = 333
  qp, rng = _harvest(s, b, 20, 23)
  base = s.joint_offsets()
  deltas = np.float32([[0, 0, 0], [0.02, -0.01, 0.01], [-0.02, 0.015, 0.0]])
  off = (base[None] + deltas[rng.integers(0, 3, b)][:, None, :] * (np.arange(base.shape[0]) % 2 == 0)[None, :, None]).astype(np.float32)
  fr = rng.choice(np.float32([0.5, 1.0, 1.5]), b).astype(np.float32)
  act = rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32)
  got = _flat(*s.step(qp, torch.from_numpy(act).cuda(), friction=torch.from_numpy(fr).cuda(),
                      joint_offset=torch.from_numpy(off).cuda()))
  ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
  check_step('humanoid', ins, act, got, oracles=(FrictionOracle(d, fr, np.float32, joint_offset=off),
                                                 FrictionOracle(d, fr, np.float64, joint_offset=off)))
  # rollout == chained steps, bitwise
  acts = torch.from_numpy(rng.uniform(-1, 1, (4, b, s.act_size)).astype(np.float32)).cuda()
  kw = dict(friction=torch.from_numpy(fr).cuda(), joint_offset=torch.from_numpy(off).cuda())
  qr, _ = s.rollout(qp, acts, **kw)
  qc = qp
  for t in range(4):
    qc, _ = s.step(qc, acts[t], **kw)
  assert torch.equal(qr.pos, qc.pos) and torch.equal(qr.ang, qc.ang)
  # the env + in-kernel policy on top (ant: the tile keeps two per SM): the policy rollout == the acting loop
  env = envs.create('ant')
  no, na = env.observation_size, env.action_size
  qpa, rng = _harvest(env.sys, 200, 5, 29)
  offa = torch.from_numpy((env.sys.joint_offsets()[None] + rng.uniform(-0.02, 0.02, (200, 1, 3)).astype(np.float32))).cuda().contiguous()
  env.randomize_joints(offa)
  g = torch.Generator(device='cuda'); g.manual_seed(1)
  params = [((torch.rand(no, 16, generator=g, device='cuda') - .5) * .3, torch.zeros(16, device='cuda')),
            ((torch.rand(16, 2 * na, generator=g, device='cuda') - .5) * .3, torch.zeros(2 * na, device='cuda'))]
  obs0 = env.sys.observe(qpa)
  q2, _, eo = env.sys.rollout(qpa, None, env=True, policy=Policy(params, activation='tanh'), obs0=obs0, steps=3, info=False,
                              joint_offset=offa)
  st = envs.State(qpa, obs0, None, None)
  for t in range(3):
    st = env.step(st, eo['action'][t].contiguous())
    assert torch.equal(st.obs, eo['obs'][t]), t
  assert torch.equal(st.qp.pos, q2.pos)


def _reset_desc_with_offsets(desc, off):
  """`desc` with the reset tables' parent offsets replaced by off[J,3] (step-joint order): reset joints are matched
  to the step's joints by their bodies, as bx_system_create does."""
  import copy
  d = copy.copy(desc)
  d.reset = copy.deepcopy(desc.reset)
  step = {(j['body_p'], j['body_c']): i for i, j in enumerate(desc.joints)}
  for rj in d.reset['joints']:
    rj['off_p'] = [float(x) for x in off[step[(rj['body_p'], rj['body_c'])]]]
  return d


@pytest.mark.parametrize('name', JOINT_SYSTEMS)
def test_reset_oracle_reproduces_reference_default_qp_under_substituted_joint_offsets(name):
  from oracle import reset_oracle
  g, _, _ = _golden_joints(name)
  desc = util.load_desc(name)
  ja, jv = g['reset_joint_angle'][None], g['reset_joint_velocity'][None]
  for v in range(g['joint_offset'].shape[0]):
    got = reset_oracle.default_qp(_reset_desc_with_offsets(desc, g['joint_offset'][v]), ja, jv, np.float64)
    for k in util.QP_KEYS:
      np.testing.assert_allclose(got[k][0], g['reset_' + k][v], rtol=0, atol=1e-12, err_msg='%s variant %d' % (k, v))
  assert np.abs(g['reset_pos'][1] - g['reset_pos'][0]).max() > 5e-3     # the sockets move the pose


@pytest.mark.gpu
@pytest.mark.parametrize('name', JOINT_SYSTEMS)
def test_gpu_default_qp_with_per_env_joint_offsets(name):
  """bx_default_qp_offsets: ONE launch over envs with different sockets == the reference's default_qp of each variant."""
  torch = _torch()
  from tests.test_gpu_parity import _system
  g, _, _ = _golden_joints(name)
  s = _system(name)
  nv = g['joint_offset'].shape[0]
  reps = 40                                   # ragged: 160 envs = 5 tiles, variants interleaved
  idx = np.arange(nv * reps) % nv
  off = torch.from_numpy(g['joint_offset'][idx].astype(np.float32)).cuda()
  ja = torch.from_numpy(np.tile(g['reset_joint_angle'].astype(np.float32), (nv * reps, 1))).cuda()
  jv = torch.from_numpy(np.tile(g['reset_joint_velocity'].astype(np.float32), (nv * reps, 1))).cuda()
  qp = s.default_qp(joint_angle=ja, joint_velocity=jv, joint_offset=off)
  for k in util.QP_KEYS:
    got = getattr(qp, k).cpu().numpy()
    np.testing.assert_allclose(got, g['reset_' + k][idx], rtol=0, atol=2e-6, err_msg=k)
  plain = s.default_qp(joint_angle=ja, joint_velocity=jv)
  assert torch.equal(plain.pos[idx == 0], qp.pos[idx == 0])          # the config's lanes: bitwise the plain launch
  assert float((plain.pos[idx == 1] - qp.pos[idx == 1]).abs().max()) > 5e-3


@pytest.mark.gpu
def test_gpu_env_reset_uses_randomised_joint_sockets():
  torch = _torch()
  from brax_b200 import envs
  env = envs.create('ant')
  b = 64
  base = torch.from_numpy(env.sys.joint_offsets()).cuda()
  off = (base[None] + torch.linspace(-0.03, 0.03, b, device='cuda')[:, None, None]).contiguous()
  st0 = env.reset(b, device='cuda', seed=5)
  env.randomize_joints(off)
  st1 = env.reset(b, device='cuda', seed=5)
  want = env.sys.default_qp(joint_angle=None, joint_velocity=torch.zeros(b, env.sys.num_joint_dof, device='cuda'), joint_offset=off)
  assert st1.qp.pos.shape == st0.qp.pos.shape and float((st1.qp.pos - st0.qp.pos).abs().max()) > 5e-3
  mid = b // 2      # linspace passes (almost) through 0 in the middle: nearly the config's pose there
  assert float((st1.qp.pos[mid] - st0.qp.pos[mid]).abs().max()) < 2e-3
  assert want.pos.shape == st1.qp.pos.shape
