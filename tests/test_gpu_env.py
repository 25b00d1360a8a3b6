"""Parity of the fused env epilogue ("next" row N1) with the oracle (GPU box only).

`brax_b200.envs.Ant / Humanoid .step` = ONE launch of the step kernel that also
writes what the reference's `Env.step` computes after `sys.step`: obs, reward,
done, metrics (brax/envs/ant.py:222-281, brax/envs/humanoid.py:246-340).  Checked
against the float32 oracle chain (C physics oracle -> oracle/env_oracle.py, pinned
to the reference by tests/test_env_oracle.py) under the same policy as the physics
parity tests: |kernel - o32| <= 1e-6 + 1e-5 |o32| + GAP_MULT * gap(env) + velocity
floor, gap = max |o32 - o64| of that env (measured float32 conditioning).
"""
import os

import numpy as np
import pytest

from tests import util
from tests.test_env_oracle import VARIANTS, env_golden
from tests.test_gpu_parity import GAP_MULT, OUTLIER_FRAC, VEL_ULPS, _harvest, _torch

pytestmark = pytest.mark.gpu

# no reference golden (the env goldens cover the pbd config): checked against the oracle
# chain only -- exercises the egocentric rows in JS (legacy_spring has no dead pose rows)
from brax_b200 import envs as _envs  # noqa: E402
VARIANTS = dict(VARIANTS, fetch_spring=('fetch_spring', _envs.fetch_desc(util.load_desc('fetch_spring'))))


def _env(variant, device=0):
  from brax_b200 import envs
  sysname, e = VARIANTS[variant]
  kw = dict(use_contact_forces=e.use_contact_forces) if e.kind == 'ant' else {}
  cls = {'ant': envs.Ant, 'humanoid': envs.Humanoid, 'fetch': envs.Fetch, 'grasp': envs.Grasp}[e.kind]
  env = cls(legacy_spring=sysname.endswith('_spring'), device=device, **kw)
  assert env.env == e  # the mirrored constructor defaults ARE the reference's
  return sysname, env


def _oracle_chain(sysname, envdesc, ins, act, dtype):
  from oracle import env_oracle
  from oracle.oracle import Oracle
  desc = util.load_desc(sysname)
  o = Oracle(desc, dtype)
  pos_in, act_in = env_oracle.env_prologue(desc, envdesc, ins[0], act, dtype)   # grasp.py:75-85
  out = o.step(pos_in, ins[1], ins[2], ins[3], act_in)
  fa, fv = env_oracle.flat_angle_vel(desc, *o.angle_vel(out['rot'], out['ang']))
  r = env_oracle.env_epilogue(desc, envdesc, ins[0], out, act, fa, fv, dtype)
  r['qp'] = {k: out[k] for k in util.QP_KEYS}
  return r


def check_env(sysname, envdesc, ins, act, got, outlier_frac=OUTLIER_FRAC):
  r32 = _oracle_chain(sysname, envdesc, ins, act, np.float32)
  r64 = _oracle_chain(sysname, envdesc, ins, act, np.float64)
  b = r32['obs'].shape[0]
  allowed = int(outlier_frac * b) if b >= 200 else 0
  floor = VEL_ULPS * 6e-8 / util.load_desc(sysname).dt_sub
  bad = np.zeros(b, bool)
  for k in ('obs', 'reward', 'metrics'):
    g, r = got[k].reshape(b, -1), r32[k].reshape(b, -1)
    assert np.isfinite(g).all(), k
    gap = np.abs(r.astype(np.float64) - r64[k].reshape(b, -1)).max(axis=1, keepdims=True)
    tol = 1e-6 + 1e-5 * np.abs(r) + GAP_MULT * gap + floor
    over = np.abs(g - r) > tol
    assert not (np.abs(g - r) > 100 * tol).any(), (k, float(np.abs(g - r).max()))
    bad |= over.any(axis=1)
  if envdesc.kind in ('ant', 'humanoid'):
    # done is a float compare of the torso height with the healthy range: exact unless
    # the height sits within rounding of a bound
    z = r64['obs'][:, 0]
    near = np.minimum(np.abs(z - envdesc.healthy_z_range[0]), np.abs(z - envdesc.healthy_z_range[1])) < 1e-5
    assert ((got['done'] == r32['done']) | near).all()
  else:
    assert (got['done'] == 0).all()
  assert bad.sum() <= allowed, ('envs outside tolerance', int(bad.sum()), b)
  return r32


@pytest.mark.parametrize('variant', [v for v in sorted(VARIANTS) if v != 'fetch_spring'])
def test_env_step_on_golden_inputs(variant):
  """Host path (numpy in, numpy out) on the reference's own stored steps; also
  close to the reference's float64 obs / reward (tests/golden/*_env.npz)."""
  _torch()
  from brax_b200.envs import State
  from brax_b200.system import QP
  sysname, env = _env(variant)
  g, e = util.load_golden(sysname), env_golden(variant)
  ins = [g[k].astype(np.float32) for k in util.QP_KEYS]
  act = g['act'].astype(np.float32)
  ego = env.env.kind in ('fetch', 'grasp')
  kw = dict(rand=e['rand'].astype(np.float32)) if ego else {}   # the reference's own draws
  st = env.step(State(QP(*ins), None, None, None), act, **kw)
  got = {'obs': st.obs, 'reward': st.reward, 'done': st.done,
         'metrics': np.stack([st.metrics[k] for k in env.env.metric_names()], -1)}
  assert st.obs.shape == e['obs'].shape and st.obs.shape[1] == env.observation_size
  r32 = check_env(sysname, env.env, ins, act, got)
  np.testing.assert_array_equal(st.done, e['done'])
  np.testing.assert_allclose(st.reward, e['reward'], rtol=1e-4, atol=5e-3)
  assert np.quantile(np.abs(st.obs - e['obs']), 0.99) < 1e-3
  if ego:
    # the step's QP is that of the oracle chain (grasp: action map + palm move inside the launch)
    # with hit targets teleported by the reference's draws, and that of the reference itself
    from oracle import env_oracle
    want = env_oracle.env_teleport(env.env, r32['qp']['pos'], r32['metrics'][:, 0], e['rand'], np.float32)
    np.testing.assert_allclose(st.qp.pos, want, rtol=1e-5, atol=1e-4)
    np.testing.assert_allclose(st.qp.pos, e['qp_pos'], rtol=1e-5, atol=2e-4)
    if env.env.kind == 'grasp':
      palm = env.env.ref_body
      assert np.abs(st.qp.pos[:, palm] - ins[0][:, palm]).max() > 1e-3   # the palm did move
  else:
    # the physics outputs of the env step are those of the plain step
    qp2, _ = env.sys.step(QP(*ins), act)
    np.testing.assert_array_equal(qp2.pos, st.qp.pos)


@pytest.mark.parametrize('variant,b,steps', [('ant', 2048, 30), ('humanoid', 2048, 30),
                                             ('ant_spring', 1024, 20), ('humanoid_spring', 1024, 20),
                                             ('fetch', 1024, 20), ('grasp', 1024, 20),
                                             ('fetch_spring', 1024, 20)])
def test_env_step_on_harvested_states(variant, b, steps):
  """Device path (torch tensors) on states with active contacts."""
  torch = _torch()
  from brax_b200.envs import State
  sysname, env = _env(variant)
  qp, rng = _harvest(env.sys, b, steps, 7)
  act = rng.uniform(-1, 1, (b, env.action_size)).astype(np.float32)
  if env.env.kind in ('fetch', 'grasp'):
    # put some targets inside the hit radius so `hits` and the 5 x hit / facing terms are live
    pos = qp.pos.clone()
    src = env.env.ref_body if env.env.kind == 'fetch' else env.env.object_body
    far = torch.tensor([8.0, 3.0, 1.0], device=pos.device)
    pos[:, env.env.target_body] = pos[:, src] + far      # as after the reference's reset: away
    pos[: b // 4, env.env.target_body] = pos[: b // 4, src] + 0.3
    qp = type(qp)(pos, qp.rot, qp.vel, qp.ang)
  kw = {}
  if env.env.kind in ('fetch', 'grasp'):
    rand = rng.uniform(0, 1, (b, 3)).astype(np.float32)
    kw = dict(rand=torch.from_numpy(rand).cuda())
  st = env.step(State(qp, None, None, None), torch.from_numpy(act).cuda(), **kw)
  torch.cuda.synchronize()
  got = {'obs': st.obs.cpu().numpy(), 'reward': st.reward.cpu().numpy(), 'done': st.done.cpu().numpy(),
         'metrics': torch.stack([st.metrics[k] for k in env.env.metric_names()], -1).cpu().numpy()}
  ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
  r32 = check_env(sysname, env.env, ins, act, got)
  if env.env.kind == 'ant':
    assert (np.abs(r32['obs'][:, 27:]) > 0).mean() > 0.02  # contact forces are live
  if env.env.kind in ('fetch', 'grasp'):
    hit = got['metrics'][:, 0] > 0
    assert 0.1 < hit.mean() < 0.9
    # hit targets were teleported inside the launch with the given uniforms, the others left alone
    from oracle import env_oracle
    want = env_oracle.env_teleport(env.env, r32['qp']['pos'], got['metrics'][:, 0], rand, np.float32)
    np.testing.assert_allclose(st.qp.pos.cpu().numpy()[:, env.env.target_body], want[:, env.env.target_body],
                               rtol=1e-5, atol=1e-4)
    # without env_rand the launch leaves every target where it is
    _, _, eo = env.sys.step(qp, torch.from_numpy(act).cuda(), env=True)
    assert torch.equal(eo['obs'], st.obs)


@pytest.mark.parametrize('variant', ['fetch', 'grasp'])
def test_obs_is_built_before_the_target_teleport(variant):
  """Every env scores a hit, so every target is teleported inside the launch; the obs (target
  body's pos_local included) must still be that of the state BEFORE the teleport (fetch.py:56
  vs :88-93, grasp.py:88 vs :131-136) for every CTA size, i.e. whichever warp owns the target
  body's rows (the teleporting warp itself: no other warp reads the row it rewrites)."""
  torch = _torch()
  from brax_b200.envs import State
  from oracle import env_oracle
  sysname, env = _env(variant)
  b = 512
  qp, rng = _harvest(env.sys, b, 10, 3)
  act = rng.uniform(-1, 1, (b, env.action_size)).astype(np.float32)
  pos = qp.pos.clone()
  src = env.env.ref_body if env.env.kind == 'fetch' else env.env.object_body
  pos[:, env.env.target_body] = pos[:, src] + 0.05
  qp = type(qp)(pos, qp.rot, qp.vel, qp.ang)
  rand = rng.uniform(0, 1, (b, 3)).astype(np.float32)
  ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
  for w in (8, 10, 12, 16):
    env.sys.configure(threads_per_block=32 * w)
    for _ in range(3):   # the race this guards against was timing dependent
      st = env.step(State(qp, None, None, None), torch.from_numpy(act).cuda(), rand=torch.from_numpy(rand).cuda())
      torch.cuda.synchronize()
      got = {'obs': st.obs.cpu().numpy(), 'reward': st.reward.cpu().numpy(), 'done': st.done.cpu().numpy(),
             'metrics': torch.stack([st.metrics[k] for k in env.env.metric_names()], -1).cpu().numpy()}
      r32 = check_env(sysname, env.env, ins, act, got)
      assert (got['metrics'][:, 0] > 0).mean() > 0.9, w
      want = env_oracle.env_teleport(env.env, r32['qp']['pos'], got['metrics'][:, 0], rand, np.float32)
      tb = env.env.target_body
      np.testing.assert_allclose(st.qp.pos.cpu().numpy()[:, tb], want[:, tb], rtol=1e-5, atol=1e-4)
      assert np.abs(want[:, tb] - r32['qp']['pos'][:, tb]).max() > 0.5   # the targets did move


@pytest.mark.parametrize('variant', ['ant', 'humanoid'])
def test_env_step_with_host_io(variant):
  """Gym-style loop with the state resident on the GPU: page-locked host actions in, obs / reward /
  done / metrics written by the launch into page-locked host tensors == the device-tensor path."""
  torch = _torch()
  from brax_b200.envs import State
  sysname, env = _env(variant)
  b = 777
  qp, rng = _harvest(env.sys, b, 5, 13)
  hio = env.host_io(b)
  st_d = st_h = State(qp, None, None, None)
  for t in range(3):
    act = torch.from_numpy(rng.uniform(-1, 1, (b, env.action_size)).astype(np.float32))
    hio['action'].copy_(act)
    st_d = env.step(st_d, act.cuda())
    st_h = env.step(st_h, hio['action'], host_out=hio)
    torch.cuda.synchronize()
    assert not st_h.obs.is_cuda and st_h.obs.data_ptr() == hio['obs'].data_ptr()
    assert torch.equal(st_h.obs, st_d.obs.cpu()) and torch.equal(st_h.reward, st_d.reward.cpu())
    assert torch.equal(st_h.done, st_d.done.cpu())
    for k in env.env.metric_names():
      assert torch.equal(st_h.metrics[k], st_d.metrics[k].cpu()), k
    assert torch.equal(st_h.qp.pos, st_d.qp.pos)
  with pytest.raises(RuntimeError, match='page-locked'):
    env.step(st_h, torch.zeros(b, env.action_size), host_out=hio)


def _mlp(torch, sizes, seed, scale=1.0):
  g = torch.Generator(device='cuda'); g.manual_seed(seed)
  return [((torch.rand(sizes[i], sizes[i + 1], generator=g, device='cuda') * 2 - 1) * scale * (3.0 / sizes[i]) ** 0.5,
           (torch.rand(sizes[i + 1], generator=g, device='cuda') * 2 - 1) * 0.1) for i in range(len(sizes) - 1)]


def _policy_torch(torch, params, act_fn, obs, mean, std, na, noise=None, min_std=0.001):
  """training/networks.py:83-102 + distribution.py:131-158 in torch float32 (the checker of the in-kernel policy)."""
  x = (obs - mean) / std
  for i, (w, b) in enumerate(params):
    x = x @ w + b
    if i < len(params) - 1:
      x = {'relu': torch.relu, 'tanh': torch.tanh, 'swish': lambda v: v * torch.sigmoid(v)}[act_fn](x)
  loc = x[:, :na]
  if noise is not None:
    loc = loc + (torch.nn.functional.softplus(x[:, na:]) + min_std) * noise
  return torch.tanh(loc)


@pytest.mark.parametrize('variant,act_fn,hidden', [('ant', 'swish', (32, 32, 32, 32)), ('humanoid', 'relu', (64, 48)),
                                                   ('humanoid', 'swish', (32, 32)),   # pre-activations below -88 (swish without overflow)
                                                   ('ant', 'tanh', (64, 64)),         # obs staged in chunks of 64 rows (two tiles per SM kept)
                                                   ('grasp', 'tanh', (40,)), ('ant', 'relu', (256, 100))])
def test_in_kernel_policy_rollout_equals_acting_loop(variant, act_fn, hidden):
  """bx_policy: T steps of actions = policy(obs); state = env.step(state, actions) in ONE launch == the
  acting loop (training/acting.py:53-74) with the policy evaluated in torch between single-step launches:
  the first step's actions to float32 rounding, the following steps within the growth of that rounding."""
  torch = _torch()
  from brax_b200.envs import State
  from brax_b200.system import Policy
  sysname, env = _env(variant)
  b, t_n = 333, 5
  na, no = env.action_size, env.observation_size
  qp, rng = _harvest(env.sys, b, 5, 51)
  if env.env.kind == 'grasp':
    pos = qp.pos.clone()
    pos[:, env.env.target_body] = pos[:, env.env.object_body] + torch.tensor([8.0, 3.0, 1.0], device='cuda')
    qp = type(qp)(pos, qp.rot, qp.vel, qp.ang)
  params = _mlp(torch, (no,) + hidden + (2 * na,), 7)
  mean = torch.from_numpy(rng.uniform(-.2, .2, no).astype(np.float32)).cuda()
  std = torch.from_numpy(rng.uniform(.5, 2., no).astype(np.float32)).cuda()
  obs0 = env.sys.observe(qp)
  for noise in (None, torch.from_numpy(rng.standard_normal((t_n, b, na)).astype(np.float32)).cuda()):
    pol = Policy(params, activation=act_fn, obs_mean=mean, obs_std=std)
    q2, _, eo = env.sys.rollout(qp, None, env=True, policy=pol, obs0=obs0, noise=noise, steps=t_n, info=False)
    torch.cuda.synchronize()
    assert eo['action'].shape == (t_n, b, na) and eo['obs'].shape == (t_n, b, no)
    st, obs = State(qp, obs0, None, None), obs0
    for t in range(t_n):
      a = _policy_torch(torch, params, act_fn, obs, mean, std, na, None if noise is None else noise[t])
      tol = 2e-5 * (4 ** t)
      assert float((eo['action'][t] - a).abs().max()) <= tol + 1e-3 * (t > 0), (variant, t, float((eo['action'][t] - a).abs().max()))
      st = env.step(st, eo['action'][t].contiguous())     # the kernel's own action: the env step itself is bitwise
      assert torch.equal(st.obs, eo['obs'][t]), (variant, t)
      assert torch.equal(st.reward, eo['reward'][t])
      obs = st.obs
    assert torch.equal(st.qp.pos, q2.pos)
  with pytest.raises(RuntimeError, match='obs size'):
    env.sys.rollout(qp, None, env=True, policy=Policy(_mlp(torch, (no + 1, 8, 2 * na), 1)), obs0=obs0, steps=2)


def test_in_kernel_policy_with_auto_reset():
  """Policy rollout under EpisodeWrapper + AutoResetWrapper in the launch: an env that was restarted acts on
  first_obs at the next step (wrappers.py:147), exactly like the acting loop around the wrapped env."""
  torch = _torch()
  from brax_b200.system import AutoReset, Policy
  sysname, env = _env('ant')
  b, t_n, length = 256, 6, 3
  na, no = env.action_size, env.observation_size
  qp, rng = _harvest(env.sys, b, 5, 53)
  params = _mlp(torch, (no, 32, 2 * na), 9)
  one = torch.ones(no, device='cuda')
  pol = Policy(params, activation='relu')
  first_qp = type(qp)(*[getattr(qp, k).clone() for k in util.QP_KEYS])
  first_obs = env.sys.observe(first_qp)
  steps0 = torch.from_numpy(rng.integers(0, length, b).astype(np.float32)).cuda()
  ar = AutoReset(first_qp, first_obs, torch.zeros(b, device='cuda'), steps0, length)
  q2, _, eo = env.sys.rollout(qp, None, env=True, policy=pol, obs0=first_obs, steps=t_n, auto_reset=ar, info=False)
  q, obs, done, steps = qp, first_obs, torch.zeros(b, device='cuda'), steps0
  for t in range(t_n):
    a = _policy_torch(torch, params, 'relu', obs, 0 * one, one, na)
    assert float((eo['action'][t] - a).abs().max()) <= 1e-4, t
    q, _, e1 = env.sys.step(q, eo['action'][t].contiguous(), env=True, info=False,
                            auto_reset=AutoReset(first_qp, first_obs, done, steps, length))
    assert torch.equal(e1['obs'], eo['obs'][t]) and torch.equal(e1['done'], eo['done'][t]), t
    obs, done, steps = e1['obs'], e1['done'], e1['steps']
  assert float(eo['done'].sum()) > 0     # episodes did end and restart inside the launch
  assert torch.equal(q.pos, q2.pos)


def test_device_state_handle_rollouts_equal_plain_rollouts():
  """bx_state (System.make_state): the QP lives in the library and is stepped in place -- two chained
  rollouts on the handle == the same rollouts through bx_rollout with the QP passed around (bitwise:
  obs / reward / done / metrics of every step, Info of the last, the final QP)."""
  torch = _torch()
  sysname, env = _env('ant')
  b, t_n = 700, 5
  qp, rng = _harvest(env.sys, b, 5, 41)
  acts = torch.from_numpy(rng.uniform(-1, 1, (2, t_n, b, env.action_size)).astype(np.float32)).cuda()
  st = env.sys.make_state(b).set(qp)
  q = qp
  for r in range(2):
    q, info_w, eo_w = env.sys.rollout(q, acts[r], env=True)
    info_g, eo_g = st.rollout(acts[r], env=True, info=True)
    for k in ('obs', 'reward', 'done', 'metrics'):
      assert torch.equal(eo_g[k], eo_w[k]), (r, k)
    assert torch.equal(info_g.contact.vel, info_w.contact.vel) and torch.equal(info_g.contact_pos, info_w.contact_pos)
  got = st.get()
  for k in util.QP_KEYS:
    assert torch.equal(getattr(got, k), getattr(q, k)), k
  # one step without env outputs, and argument errors
  st.rollout(acts[0, 0])
  with pytest.raises(ValueError):
    st.rollout(acts[0, :, :10])
  with pytest.raises(ValueError):
    st.set(type(qp)(qp.pos[:10], qp.rot[:10], qp.vel[:10], qp.ang[:10]))


def test_env_rollout_equals_repeated_steps():
  """bx_rollout writes the env outputs of EVERY step: [T,B,...] bit-identical to T
  env steps; ragged and unbatched shapes equal the same envs inside a big batch."""
  torch = _torch()
  from brax_b200.envs import State
  from brax_b200.system import QP
  _, env = _env('ant')
  b, t_n = 300, 6   # ragged: 300 = 9 tiles + 12 envs; rows of step t start unaligned
  qp, rng = _harvest(env.sys, b, 3, 3)
  acts = torch.from_numpy(rng.uniform(-1, 1, (t_n, b, env.action_size)).astype(np.float32)).cuda()
  st, obs, rew, done, met = State(qp, None, None, None), [], [], [], []
  for t in range(t_n):
    st = env.step(st, acts[t])
    obs.append(st.obs); rew.append(st.reward); done.append(st.done)
    met.append(torch.stack([st.metrics[k] for k in env.env.metric_names()], -1))
  q2, i2, eo = env.sys.rollout(qp, acts, env=True)
  torch.cuda.synchronize()
  assert torch.equal(q2.pos, st.qp.pos) and torch.equal(q2.ang, st.qp.ang)
  assert torch.equal(eo['obs'], torch.stack(obs)) and torch.equal(eo['reward'], torch.stack(rew))
  assert torch.equal(eo['done'], torch.stack(done)) and torch.equal(eo['metrics'], torch.stack(met))
  # unbatched call == env 5 of the batch
  one = env.step(State(QP(*[getattr(qp, k)[5] for k in util.QP_KEYS]), None, None, None), acts[0, 5])
  assert one.obs.shape == (env.observation_size,) and one.reward.shape == ()
  assert torch.equal(one.obs, obs[0][5]) and torch.equal(one.reward, rew[0][5])


def test_env_rollout_with_teleports_and_host_buffers():
  """grasp: a rollout with per-step uniforms [T,B,3] == the same steps one by one (action map,
  palm move and teleports inside the launch every step); ant: the page-locked host path
  (zero-copy launch on HostBuffers) returns the env outputs of the device path."""
  torch = _torch()
  from brax_b200.envs import State
  from brax_b200.system import QP
  _, env = _env('grasp')
  b, t_n = 200, 5
  qp, rng = _harvest(env.sys, b, 3, 5)
  pos = qp.pos.clone()
  pos[:, env.env.target_body] = pos[:, env.env.object_body] + 0.2     # hits from the first step on
  qp = QP(pos, qp.rot, qp.vel, qp.ang)
  acts = torch.from_numpy(rng.uniform(-1, 1, (t_n, b, env.action_size)).astype(np.float32)).cuda()
  rands = torch.from_numpy(rng.uniform(0, 1, (t_n, b, 3)).astype(np.float32)).cuda()
  st, obs, rew = State(qp, None, None, None), [], []
  for t in range(t_n):
    st = env.step(st, acts[t], rand=rands[t])
    obs.append(st.obs); rew.append(st.reward)
  q2, _, eo = env.sys.rollout(qp, acts, env=True, env_rand=rands)
  torch.cuda.synchronize()
  assert torch.equal(q2.pos, st.qp.pos) and torch.equal(eo['obs'], torch.stack(obs))
  assert torch.equal(eo['reward'], torch.stack(rew)) and float(eo['metrics'][..., 0].sum()) > 0
  # host path on page-locked buffers
  _, ant = _env('ant')
  b = 96
  qpa, rng = _harvest(ant.sys, b, 3, 9)
  act = rng.uniform(-1, 1, (b, ant.action_size)).astype(np.float32)
  dq, _, deo = ant.sys.step(qpa, torch.from_numpy(act).cuda(), env=True)
  hb = ant.sys.host_buffers(b)
  for k in util.QP_KEYS:
    getattr(hb, k)[...] = getattr(qpa, k).cpu().numpy()
  hb.act[...] = act
  hq, _, heo = ant.sys.step(hb.qp(), hb.act, out=hb, env=True)
  np.testing.assert_array_equal(hq.pos, dq.pos.cpu().numpy())
  for k in ('obs', 'reward', 'done', 'metrics'):
    np.testing.assert_array_equal(heo[k], deo[k].cpu().numpy(), err_msg=k)


def test_env_outputs_need_an_env_system():
  _torch()
  from brax_b200.system import load_system
  s = load_system('ant')
  with pytest.raises(ValueError):
    s.step(s.default_qp(batch=4), np.zeros((4, 8), np.float32), env=True)


@pytest.mark.parametrize('variant,b', [('ant', 1024), ('humanoid', 1024), ('grasp', 512), ('fetch', 512)])
def test_1000_step_statistics(variant, b):
  """North star: 'reward/energy statistics over 1000-step rollouts are compared
  within a stated tolerance'.  1000 env steps of B envs on the GPU (10 launches of
  100 steps through bx_rollout, env epilogue every step) against the float32 oracle
  chain run over the same actions from the same start.  Single trajectories are
  chaotic (contact-rich PBD amplifies ulps), so the POPULATION is compared: mean
  reward per step, mean torso height, mean kinetic energy and the fraction of
  unhealthy env-steps over all B x 1000 samples agree within 2 % of the oracle's
  value (or 3 standard errors of the mean across envs, whichever is larger), and the
  first 10 steps agree per env (rollout tolerance of test_rollout_10_steps).
  grasp / fetch (BASELINE config 4 and the box-collider env): the same with the in-launch
  action map / palm move and target teleports (shared uniforms); 'height' is obs[0] (distance
  to the object / torso_fwd.x) and 'unhealthy' the hit rate."""
  torch = _torch()
  from oracle import env_oracle
  from oracle.oracle import Oracle
  sysname, env = _env(variant)
  desc = util.load_desc(sysname)
  t_n, chunk = 1000, 100
  rng = np.random.default_rng(2024)
  st0 = env.reset(batch=b, rng=rng, velocity_noise=0.1)
  acts = rng.uniform(-1, 1, (t_n, b, env.action_size)).astype(np.float32)
  mass = np.asarray([x['mass'] for x in desc.bodies], np.float32)
  live = np.asarray([x['pos_mask'][2] for x in desc.bodies], np.float32)  # skip the frozen floor

  def energy(vel):
    return 0.5 * np.sum(mass[None, :, None] * live[None, :, None] * np.square(vel), axis=(-1, -2))

  # GPU: 10 launches x 100 steps
  from brax_b200.system import QP
  qp = QP(*[torch.from_numpy(np.ascontiguousarray(getattr(st0.qp, k))).cuda() for k in util.QP_KEYS])
  g_rew, g_z, g_e, g_done, g_first = [], [], [], [], None
  ego = env.env.kind in ('fetch', 'grasp')
  tz = 0 if (ego or env.env.exclude_current_positions) else 2
  rands = rng.uniform(0, 1, (t_n, b, 3)).astype(np.float32) if ego else None
  if ego:   # targets start away from the robot, as after the reference's reset
    src = env.env.object_body if env.env.kind == 'grasp' else env.env.ref_body
    st0.qp.pos[:, env.env.target_body] = st0.qp.pos[:, src] + np.asarray([3.0, 1.0, 1.0], np.float32)
    qp = QP(*[torch.from_numpy(np.ascontiguousarray(getattr(st0.qp, k))).cuda() for k in util.QP_KEYS])
  for c0 in range(0, t_n, chunk):
    a = torch.from_numpy(acts[c0:c0 + chunk]).cuda()
    # energy needs the velocities of every step: step-by-step inside the chunk would be
    # 1000 launches; instead take z / reward / done from the per-step env outputs and
    # the energy at the chunk ends
    qp, _, eo = env.sys.rollout(qp, a, env=True,
                                env_rand=torch.from_numpy(rands[c0:c0 + chunk]).cuda() if ego else None)
    g_rew.append(eo['reward'].cpu().numpy()); g_z.append(eo['obs'][..., tz].cpu().numpy())
    g_done.append((eo['metrics'][..., 0] if ego else eo['done']).cpu().numpy())
    g_e.append(energy(qp.vel.cpu().numpy()))
    if g_first is None:
      g_first = eo['obs'][9].cpu().numpy()
  g_rew, g_z, g_done = np.concatenate(g_rew), np.concatenate(g_z), np.concatenate(g_done)
  assert np.isfinite(g_rew).all() and np.isfinite(g_z).all()

  # oracle: the same 1000 steps on the host cores
  o = Oracle(desc, np.float32)
  cur = {k: np.ascontiguousarray(getattr(st0.qp, k), np.float32) for k in util.QP_KEYS}
  o_rew, o_z, o_done, o_e, o_first = np.empty((t_n, b), np.float32), np.empty((t_n, b), np.float32), \
      np.empty((t_n, b), np.float32), [], None
  for t in range(t_n):
    pos_in, act_in = env_oracle.env_prologue(desc, env.env, cur['pos'], acts[t], np.float32)
    out = o.step(pos_in, cur['rot'], cur['vel'], cur['ang'], act_in, num_threads=os.cpu_count() or 1)
    fa, fv = env_oracle.flat_angle_vel(desc, *o.angle_vel(out['rot'], out['ang']))
    r = env_oracle.env_epilogue(desc, env.env, cur['pos'], out, acts[t], fa, fv, np.float32)
    if ego:
      out['pos'] = env_oracle.env_teleport(env.env, out['pos'], r['metrics'][:, 0], rands[t], np.float32)
    o_rew[t], o_z[t], o_done[t] = r['reward'], r['obs'][:, tz], (r['metrics'][:, 0] if ego else r['done'])
    if t == 9:
      o_first = r['obs']
    if (t + 1) % chunk == 0:
      o_e.append(energy(out['vel']))
    cur = {k: out[k] for k in util.QP_KEYS}

  # the first 10 steps agree per env
  err = np.abs(g_first - o_first).max(axis=1)
  assert np.quantile(err, 0.99) <= 5e-2 and np.median(err) <= 2e-3, (float(np.quantile(err, 0.99)), float(np.median(err)))

  def close(name, g, r):
    # per-env means -> population mean and its standard error
    gm, rm = g.mean(axis=0), r.mean(axis=0)
    se = np.sqrt(gm.var() / b + rm.var() / b)
    tol = max(0.02 * abs(rm.mean()), 3.0 * se)
    assert abs(gm.mean() - rm.mean()) <= tol, (variant, name, float(gm.mean()), float(rm.mean()), float(tol))

  close('reward', g_rew, o_rew)
  close('torso_z', g_z, o_z)
  close('unhealthy', g_done, o_done)
  close('kinetic_energy', np.stack(g_e), np.stack(o_e))


def _auto_reset_case(variant, b, seed, episode_length):
  """Harvested states, a share of them lifted out of the healthy range (ant / humanoid: done),
  random step counters / done flags, per-env first states and first obs."""
  torch = _torch()
  from brax_b200.system import QP, AutoReset
  sysname, env = _env(variant)
  qp, rng = _harvest(env.sys, b, 12, seed)
  pos = qp.pos.clone()
  pos[: b // 5, :, 2] += 1.5                                   # unhealthy: above healthy_z_range
  qp = QP(pos, qp.rot, qp.vel, qp.ang)
  first, _ = _harvest(env.sys, b, 2, seed + 1)
  first_obs = torch.from_numpy(rng.normal(size=(b, env.observation_size)).astype(np.float32)).cuda()
  done_in = torch.from_numpy((rng.uniform(size=b) < 0.2).astype(np.float32)).cuda()
  steps_in = torch.from_numpy(rng.integers(0, episode_length, b).astype(np.float32)).cuda()
  return sysname, env, qp, rng, AutoReset(first, first_obs, done_in, steps_in, episode_length)


@pytest.mark.parametrize('variant', ['ant', 'humanoid_spring', 'fetch'])
def test_auto_reset_matches_oracle(variant):
  """EpisodeWrapper + AutoResetWrapper inside the launch (wrappers.py:105-148) against the
  oracle's wrapped step (pinned to the reference's wrapper classes, test_env_oracle) on the
  same inputs: step counter, truncation and done exact; reset envs carry first_qp / first_obs
  bit for bit; the others stay within the env tolerance."""
  torch = _torch()
  from oracle import env_oracle
  from oracle.oracle import Oracle
  b, length = 700, 3
  sysname, env, qp, rng, ar = _auto_reset_case(variant, b, 11, length)
  act = rng.uniform(-1, 1, (b, env.action_size)).astype(np.float32)
  kw, rand = {}, None
  if env.env.kind == 'fetch':
    pos = qp.pos.clone()
    pos[::3, env.env.target_body] = pos[::3, env.env.ref_body] + 0.3      # some hits -> teleports
    qp = type(qp)(pos, qp.rot, qp.vel, qp.ang)
    rand = rng.uniform(0, 1, (b, 3)).astype(np.float32)
    kw = dict(env_rand=torch.from_numpy(rand).cuda())
  q2, _, eo = env.sys.step(qp, torch.from_numpy(act).cuda(), env=True, auto_reset=ar, **kw)
  torch.cuda.synchronize()
  got = {k: v.cpu().numpy() for k, v in eo.items()}
  desc = util.load_desc(sysname)
  ins = {k: getattr(qp, k).cpu().numpy() for k in util.QP_KEYS}
  first = {k: getattr(ar.first_qp, k).cpu().numpy() for k in util.QP_KEYS}
  want = env_oracle.wrapped_step(desc, env.env, Oracle(desc, np.float32), ins, ar.done.cpu().numpy(),
                                 ar.steps.cpu().numpy(), act, first, ar.first_obs.cpu().numpy(), length,
                                 np.float32, rand)
  np.testing.assert_array_equal(got['steps'], want['steps'])
  over = want['steps'] >= length
  assert over.any() and (~over).any()
  # the env's own done, recovered from the wrapped outputs, goes through the env check below
  inner_done = np.where(over, 1 - got['truncation'], got['done'])
  np.testing.assert_array_equal(got['done'], np.where(over, 1, inner_done))
  np.testing.assert_array_equal(got['truncation'], np.where(over, 1 - inner_done, 0))
  d = got['done'] > 0
  assert 0.2 < d.mean() < 0.9
  if env.env.kind != 'fetch':
    assert (d & ~over).any()                                  # done by the healthy range
  for k in util.QP_KEYS:                                      # wrappers.py:146
    np.testing.assert_array_equal(getattr(q2, k).cpu().numpy()[d], first[k][d], err_msg=k)
  np.testing.assert_array_equal(got['obs'][d], ar.first_obs.cpu().numpy()[d])   # wrappers.py:147
  inner = _oracle_chain(sysname, env.env, [ins[k] for k in util.QP_KEYS], act, np.float32)
  obs = np.where(d[:, None], inner['obs'], got['obs'])
  check_env(sysname, env.env, [ins[k] for k in util.QP_KEYS], act,
            dict(obs=obs, reward=got['reward'], done=inner_done, metrics=got['metrics']))
  keep = ~d
  for k in util.QP_KEYS:
    w = want['qp'][k][keep]
    if env.env.kind == 'fetch' and k == 'pos':               # teleported targets: checked elsewhere
      w, g = np.delete(w, env.env.target_body, 1), np.delete(getattr(q2, k).cpu().numpy()[keep], env.env.target_body, 1)
    else:
      g = getattr(q2, k).cpu().numpy()[keep]
    frac = (np.abs(g - w) > 1e-4 + 1e-3 * np.abs(w)).reshape(len(w), -1).any(1).mean()
    assert frac <= OUTLIER_FRAC, (k, frac)


@pytest.mark.parametrize('variant', ['ant', 'humanoid', 'grasp', 'fetch_spring'])
def test_auto_reset_rollout_equals_repeated_steps(variant):
  """The (done, steps) chain of a rollout stays inside the kernel: T wrapped steps in one launch
  are bit-identical to T single wrapped launches fed with each other's done / steps."""
  torch = _torch()
  import dataclasses
  b, t_n, length = 300, 7, 3
  _, env, qp, rng, ar = _auto_reset_case(variant, b, 5, length)
  acts = torch.from_numpy(rng.uniform(-1, 1, (t_n, b, env.action_size)).astype(np.float32)).cuda()
  kw = lambda t: {}  # noqa: E731
  rand = None
  if env.env.kind in ('fetch', 'grasp'):
    rand = torch.from_numpy(rng.uniform(0, 1, (t_n, b, 3)).astype(np.float32)).cuda()
    kw = lambda t: dict(env_rand=rand[t])  # noqa: E731
  cur, a, seq = qp, ar, []
  for t in range(t_n):
    cur, _, eo = env.sys.step(cur, acts[t], env=True, auto_reset=a, **kw(t))
    seq.append(eo)
    a = dataclasses.replace(a, done=eo['done'], steps=eo['steps'])
  q2, _, er = env.sys.rollout(qp, acts, env=True, auto_reset=ar, env_rand=rand)
  torch.cuda.synchronize()
  for k in util.QP_KEYS:
    assert torch.equal(getattr(q2, k), getattr(cur, k)), k
  for k in ('obs', 'reward', 'done', 'metrics', 'truncation'):
    assert torch.equal(er[k], torch.stack([s[k] for s in seq])), k
  assert torch.equal(er['steps'], seq[-1]['steps'])
  done = er['done'].cpu().numpy()
  assert done[1:].any() and er['truncation'].any() and float(er['steps'].max()) <= length


def test_auto_reset_on_reference_sequences():
  """tests/golden/ant_wrapped_env.npz (the reference's AutoResetWrapper(EpisodeWrapper(Ant)),
  oracle/gen_golden_wrappers.py): each stored step through one wrapped launch, and each whole
  sequence through the wrapper mirror `envs.create('ant', episode_length, auto_reset=True)`."""
  torch = _torch()
  from brax_b200 import envs
  from brax_b200.system import QP, AutoReset
  g = np.load(os.path.join(util.GOLDEN, 'ant_wrapped_env.npz'))
  length = int(g['episode_length'])
  env = envs.create('ant', episode_length=length, auto_reset=True)
  assert isinstance(env, envs.EpisodeAutoReset)
  cu = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()  # noqa: E731
  first = QP(*[cu(g['first_' + k]) for k in util.QP_KEYS])
  nseq, t_n = g['done'].shape
  for t in range(t_n):
    ar = AutoReset(first, cu(g['first_obs']), cu(g['done_in'][:, t]), cu(g['steps_in'][:, t]), length)
    q2, _, eo = env.sys.step(QP(*[cu(g[k][:, t]) for k in util.QP_KEYS]), cu(g['act'][:, t]), env=True, auto_reset=ar)
    for k in ('done', 'steps', 'truncation'):
      np.testing.assert_array_equal(eo[k].cpu().numpy(), g[k][:, t], err_msg='%s step %d' % (k, t))
    np.testing.assert_allclose(eo['obs'].cpu().numpy(), g['obs'][:, t], rtol=1e-3, atol=2e-3)
    np.testing.assert_allclose(eo['reward'].cpu().numpy(), g['reward'][:, t], rtol=1e-3, atol=2e-3)
    for k in util.QP_KEYS:
      np.testing.assert_allclose(getattr(q2, k).cpu().numpy(), g[k + '_o'][:, t], rtol=1e-3, atol=2e-3, err_msg=k)
  # the chained sequence through the wrapper mirror (reset -> state.info, then step / rollout)
  st = envs.State(QP(*[cu(g[k][:, 0]) for k in util.QP_KEYS]), cu(g['first_obs']), cu(np.zeros(nseq)),
                  cu(g['done_in'][:, 0]), {}, dict(first_qp=first, first_obs=cu(g['first_obs']), steps=cu(g['steps_in'][:, 0])))
  last, eo = env.rollout(st, cu(np.swapaxes(g['act'], 0, 1)))
  for k in ('done', 'truncation'):
    np.testing.assert_array_equal(eo[k].cpu().numpy().T, g[k], err_msg=k)
  np.testing.assert_array_equal(last.info['steps'].cpu().numpy(), g['steps'][:, -1])
  np.testing.assert_allclose(np.swapaxes(eo['obs'].cpu().numpy(), 0, 1), g['obs'], rtol=5e-3, atol=1e-2)
  for t in range(t_n):
    st = env.step(st, cu(g['act'][:, t]))
    np.testing.assert_array_equal(st.done.cpu().numpy(), g['done'][:, t])
    assert torch.equal(st.obs, eo['obs'][t])


def test_auto_reset_argument_errors():
  torch = _torch()
  from brax_b200.system import AutoReset
  _, env = _env('ant')
  st = env.reset(4, device='cuda')
  act = torch.zeros(4, env.action_size, device='cuda')
  ar = AutoReset(st.qp, st.obs, st.done, None, 5)
  with pytest.raises(ValueError):
    env.sys.step(st.qp, act, auto_reset=ar)                    # needs env=True
  _, _, eo = env.sys.step(st.qp, act, env=True, auto_reset=ar)  # steps None: zeros
  assert (eo['steps'] == 1).all() and (eo['truncation'] == 0).all()
  from brax_b200 import system as bsys
  plain = bsys.load_system('ant')
  with pytest.raises(ValueError):
    plain.step(st.qp, act, env=True, auto_reset=ar)


def test_env_step_without_info_outputs():
  """info=False leaves the Info stores out of the launch (the env mirror's default, as the
  reference's Env.step drops the Info): same QP and env outputs, info is None."""
  torch = _torch()
  from brax_b200.envs import State
  _, env = _env('ant')
  qp, rng = _harvest(env.sys, 300, 5, 2)
  act = torch.from_numpy(rng.uniform(-1, 1, (300, env.action_size)).astype(np.float32)).cuda()
  q1, i1, e1 = env.sys.step(qp, act, env=True)
  q2, i2, e2 = env.sys.step(qp, act, env=True, info=False)
  q3, i3 = env.sys.step(qp, act, info=False)
  assert i1 is not None and i2 is None and i3 is None
  for k in util.QP_KEYS:
    assert torch.equal(getattr(q1, k), getattr(q2, k)) and torch.equal(getattr(q1, k), getattr(q3, k))
  for k in e1:
    assert torch.equal(e1[k], e2[k]), k
  st = env.step(State(qp, None, None, None), act)
  assert 'physics' not in st.info and torch.equal(st.obs, e1['obs'])
  env.keep_info = True
  st = env.step(State(qp, None, None, None), act)
  assert torch.equal(st.info['physics'].contact.vel, i1.contact.vel)


def test_eval_wrapper_on_reference_sequences():
  """envs.create('ant', episode_length, auto_reset=True, eval_metrics=True) against the reference's
  EvalWrapper(AutoResetWrapper(EpisodeWrapper(Ant))) sequences (ant_wrapped_env.npz): stepwise and
  as one rollout launch."""
  torch = _torch()
  from brax_b200 import envs
  from brax_b200.system import QP
  g = np.load(os.path.join(util.GOLDEN, 'ant_wrapped_env.npz'))
  length = int(g['episode_length'])
  env = envs.create('ant', episode_length=length, auto_reset=True, eval_metrics=True)
  assert isinstance(env, envs.EvalWrapper)
  cu = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()  # noqa: E731
  names = [str(k) for k in g['metric_names']]
  nseq, t_n = g['done'].shape

  def start():
    first = QP(*[cu(g['first_' + k]) for k in util.QP_KEYS])
    zero = cu(np.zeros(nseq))
    st = envs.State(QP(*[cu(g[k][:, 0]) for k in util.QP_KEYS]), cu(g['first_obs']), zero, cu(g['done_in'][:, 0]),
                    {k: zero for k in names + ['reward']},
                    dict(first_qp=first, first_obs=cu(g['first_obs']), steps=cu(g['steps_in'][:, 0])))
    st.info['eval_metrics'] = envs.EvalMetrics({k: torch.zeros_like(zero) for k in names + ['reward']},
                                               torch.ones_like(zero), torch.zeros_like(zero))
    return st

  st = start()
  for t in range(t_n):
    st = env.step(st, cu(g['act'][:, t]))
    em = st.info['eval_metrics']
    np.testing.assert_array_equal(em.active_episodes.cpu().numpy(), g['eval_active_episodes'][:, t])
    np.testing.assert_array_equal(em.episode_steps.cpu().numpy(), g['eval_episode_steps'][:, t])
    got = torch.stack([em.episode_metrics[k] for k in names + ['reward']], -1).cpu().numpy()
    np.testing.assert_allclose(got, g['eval_episode_metrics'][:, t], rtol=2e-3, atol=5e-3)
  last, _ = env.rollout(start(), cu(np.swapaxes(g['act'], 0, 1)))
  e1, e2 = st.info['eval_metrics'], last.info['eval_metrics']
  assert torch.equal(e1.active_episodes, e2.active_episodes) and torch.equal(e1.episode_steps, e2.episode_steps)
  for k in names + ['reward']:
    assert torch.allclose(e1.episode_metrics[k], e2.episode_metrics[k], rtol=1e-5, atol=1e-5), k
  # from the mirror's own reset: a 1000-step-limit eval of random actions stays consistent
  ev = envs.create('ant', episode_length=6, auto_reset=True, eval_metrics=True)
  s0 = ev.reset(128, device='cuda', seed=4)
  acts = torch.rand(9, 128, ev.action_size, device='cuda') * 2 - 1
  a, _ = ev.rollout(s0, acts)
  b = s0
  for t in range(9):
    b = ev.step(b, acts[t])
  ea, eb = a.info['eval_metrics'], b.info['eval_metrics']
  assert float(ea.active_episodes.sum()) == 0 and float(ea.episode_steps.max()) == 6
  assert torch.equal(ea.episode_steps, eb.episode_steps)
  assert torch.allclose(ea.episode_metrics['reward'], eb.episode_metrics['reward'], rtol=1e-5, atol=1e-5)


def test_eval_wrapper_rollout_with_action_repeat():
  """EvalWrapper.rollout == chained EvalWrapper.step with action_repeat = 2: `steps` advances by
  action_repeat per wrapped step (wrappers.py:112), so does eval_metrics.episode_steps."""
  torch = _torch()
  from brax_b200 import envs
  ev = envs.create('ant', episode_length=8, action_repeat=2, auto_reset=True, eval_metrics=True)
  s0 = ev.reset(96, device='cuda', seed=9)
  acts = torch.rand(6, 96, ev.action_size, device='cuda') * 2 - 1
  a, _ = ev.rollout(s0, acts)
  b = s0
  for t in range(6):
    b = ev.step(b, acts[t])
  ea, eb = a.info['eval_metrics'], b.info['eval_metrics']
  assert float(eb.episode_steps.max()) == 8 and float(ea.active_episodes.sum()) == 0
  assert torch.equal(ea.episode_steps, eb.episode_steps)
  assert torch.equal(ea.active_episodes, eb.active_episodes)
  assert torch.allclose(ea.episode_metrics['reward'], eb.episode_metrics['reward'], rtol=1e-5, atol=1e-5)


def test_action_repeat_on_reference_sequences():
  """EpisodeWrapper(action_repeat=2) inside the launch against the reference's wrapper stack
  (ant_wrapped_repeat_env.npz): two env steps per wrapped step on one action, summed rewards,
  steps += 2; stepwise, as one rollout launch, and through the create() mirror."""
  torch = _torch()
  from brax_b200 import envs
  from brax_b200.system import QP, AutoReset
  g = np.load(os.path.join(util.GOLDEN, 'ant_wrapped_repeat_env.npz'))
  length, rep = int(g['episode_length']), int(g['action_repeat'])
  assert rep == 2
  env = envs.create('ant', episode_length=length, action_repeat=rep, auto_reset=True)
  cu = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()  # noqa: E731
  first = QP(*[cu(g['first_' + k]) for k in util.QP_KEYS])
  nseq, t_n = g['done'].shape
  for t in range(t_n):
    ar = AutoReset(first, cu(g['first_obs']), cu(g['done_in'][:, t]), cu(g['steps_in'][:, t]), length, rep)
    q2, _, eo = env.sys.step(QP(*[cu(g[k][:, t]) for k in util.QP_KEYS]), cu(g['act'][:, t]), env=True, auto_reset=ar)
    for k in ('done', 'steps', 'truncation'):
      np.testing.assert_array_equal(eo[k].cpu().numpy(), g[k][:, t], err_msg='%s step %d' % (k, t))
    np.testing.assert_allclose(eo['reward'].cpu().numpy(), g['reward'][:, t], rtol=2e-3, atol=4e-3)
    np.testing.assert_allclose(eo['obs'].cpu().numpy(), g['obs'][:, t], rtol=2e-3, atol=4e-3)
    for k in util.QP_KEYS:
      np.testing.assert_allclose(getattr(q2, k).cpu().numpy(), g[k + '_o'][:, t], rtol=2e-3, atol=4e-3, err_msg=k)
  st = envs.State(QP(*[cu(g[k][:, 0]) for k in util.QP_KEYS]), cu(g['first_obs']), cu(np.zeros(nseq)),
                  cu(g['done_in'][:, 0]), {}, dict(first_qp=first, first_obs=cu(g['first_obs']), steps=cu(g['steps_in'][:, 0])))
  last, eo = env.rollout(st, cu(np.swapaxes(g['act'], 0, 1)))
  for k in ('done', 'truncation'):
    np.testing.assert_array_equal(eo[k].cpu().numpy().T, g[k], err_msg=k)
  np.testing.assert_array_equal(last.info['steps'].cpu().numpy(), g['steps'][:, -1])
  np.testing.assert_allclose(eo['reward'].cpu().numpy().T, g['reward'], rtol=5e-3, atol=1e-2)


@pytest.mark.parametrize('variant', ['ant', 'fetch_spring'])
def test_action_repeat_matches_oracle_and_chained_launches(variant):
  torch = _torch()
  import dataclasses
  from oracle import env_oracle
  from oracle.oracle import Oracle
  b, length, rep, t_n = 300, 7, 3, 4
  sysname, env, qp, rng, ar = _auto_reset_case(variant, b, 9, length)
  ar = dataclasses.replace(ar, action_repeat=rep)
  acts = torch.from_numpy(rng.uniform(-1, 1, (t_n, b, env.action_size)).astype(np.float32)).cuda()
  rand = None
  if env.env.kind == 'fetch':
    rand = torch.from_numpy(rng.uniform(0, 1, (t_n * rep, b, 3)).astype(np.float32)).cuda()
  kw = lambda t: {} if rand is None else dict(env_rand=rand[t * rep:(t + 1) * rep])  # noqa: E731
  # one wrapped step against the float32 oracle
  q1, _, e1 = env.sys.step(qp, acts[0], env=True, auto_reset=ar, **kw(0))
  desc = util.load_desc(sysname)
  ins = {k: getattr(qp, k).cpu().numpy() for k in util.QP_KEYS}
  first = {k: getattr(ar.first_qp, k).cpu().numpy() for k in util.QP_KEYS}
  want = env_oracle.wrapped_step(desc, env.env, Oracle(desc, np.float32), ins, ar.done.cpu().numpy(),
                                 ar.steps.cpu().numpy(), acts[0].cpu().numpy(), first, ar.first_obs.cpu().numpy(),
                                 length, np.float32, None if rand is None else rand[:rep].cpu().numpy(), rep)
  np.testing.assert_array_equal(e1['steps'].cpu().numpy(), want['steps'])
  got_done = e1['done'].cpu().numpy()
  assert (got_done == want['done']).mean() > 0.99
  same = got_done == want['done']
  frac = (np.abs(e1['reward'].cpu().numpy() - want['reward'])[same] > 1e-3 + 1e-3 * np.abs(want['reward'][same])).mean()
  assert frac <= 0.02, frac
  d = got_done > 0
  np.testing.assert_array_equal(q1.pos.cpu().numpy()[d], first['pos'][d])
  # T wrapped steps in one launch == T chained wrapped launches, bit for bit
  cur, a, seq = qp, ar, []
  for t in range(t_n):
    cur, _, eo = env.sys.step(cur, acts[t], env=True, auto_reset=a, **kw(t))
    seq.append(eo)
    a = dataclasses.replace(a, done=eo['done'], steps=eo['steps'])
  q2, _, er = env.sys.rollout(qp, acts, env=True, auto_reset=ar, env_rand=rand)
  torch.cuda.synchronize()
  for k in util.QP_KEYS:
    assert torch.equal(getattr(q2, k), getattr(cur, k)), k
  for k in ('obs', 'reward', 'done', 'metrics', 'truncation'):
    assert torch.equal(er[k], torch.stack([s[k] for s in seq])), k
  assert er['truncation'].any() and float(er['steps'].max()) <= length + rep - 1
  # action_repeat without the wrappers' inputs, and for grasp, is refused
  with pytest.raises(RuntimeError):
    _, genv = _env('grasp')
    gq, _ = _harvest(genv.sys, 64, 2, 1)
    z = torch.zeros(64, device='cuda')
    genv.sys.step(gq, torch.zeros(64, genv.action_size, device='cuda'), env=True,
                  auto_reset=type(ar)(gq, torch.zeros(64, genv.observation_size, device='cuda'), z, z, 5, 2))
