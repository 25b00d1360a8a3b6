"""Pins the env-epilogue oracle (oracle/env_oracle.py) against the reference.

tests/golden/<variant>_env.npz hold obs / reward / done / metrics of the
UNMODIFIED reference `Ant.step` / `Humanoid.step` (brax/envs/ant.py:222-281,
brax/envs/humanoid.py:246-340) on numpy float64 for the very steps stored in
tests/golden/<system>_steps.npz (oracle/gen_golden_env.py).  Fed with the stored
physics outputs, the float64 evaluation of the oracle must reproduce them to
rounding; chained behind the C oracle (physics + angle_vel) it must too.
"""
import os

import numpy as np
import pytest

from brax_b200 import envs
from oracle import env_oracle
from oracle.oracle import Oracle
from tests import util

# variant -> (system golden, EnvDesc as the reference constructs it)
VARIANTS = {
    'ant': ('ant', envs.EnvDesc('ant', 1.0, 0.5, 5e-4, 1.0, True, (0.2, 1.0), True, True)),
    'ant_nocfrc': ('ant', envs.EnvDesc('ant', 1.0, 0.5, 5e-4, 1.0, True, (0.2, 1.0), False, True)),
    'ant_spring': ('ant_spring', envs.EnvDesc('ant', 1.0, 0.5, 5e-4, 1.0, True, (0.2, 1.0), True, True)),
    'humanoid': ('humanoid', envs.EnvDesc('humanoid', 1.25, 0.1, 0.0, 5.0, True, (0.8, 2.1), False, True)),
    'humanoid_spring': ('humanoid_spring',
                        envs.EnvDesc('humanoid', 1.25, 0.1, 0.0, 5.0, True, (0.8, 2.1), False, True)),
    'fetch': ('fetch', envs.fetch_desc(util.load_desc('fetch'))),
    'grasp': ('grasp', envs.grasp_desc(util.load_desc('grasp'))),
}


def env_golden(variant):
  return np.load(os.path.join(util.GOLDEN, variant + '_env.npz'))


@pytest.mark.parametrize('variant', [v for v in sorted(VARIANTS) if not v.startswith('grasp')])
def test_env_oracle_f64_matches_reference(variant):
  sysname, env = VARIANTS[variant]
  g, e = util.load_golden(sysname), env_golden(variant)
  av = np.load(os.path.join(util.GOLDEN, sysname + '_angle_vel.npz'))
  desc = util.load_desc(sysname)
  assert list(e['metric_names']) == list(env.metric_names())
  assert e['obs'].shape[1] == env.obs_size(desc)
  assert abs(float(e['dt']) - desc.dt) < 1e-15
  out = {k: g[k + '_o'] for k in util.QP_KEYS}
  out.update(contact_vel=g['contact_vel'], contact_ang=g['contact_ang'])
  r = env_oracle.env_epilogue(desc, env, g['pos'], out, g['act'], av['flat_angle'], av['flat_vel'], np.float64)
  for k in ('obs', 'reward', 'done', 'metrics'):
    np.testing.assert_allclose(r[k], e[k], rtol=0, atol=1e-10, err_msg=k)


@pytest.mark.parametrize('variant', sorted(VARIANTS))
def test_env_oracle_chain_matches_reference(variant):
  """C oracle (step + angle_vel) -> env oracle, float64 end to end, and the float32
  chain (the CUDA path's checker) within float32 rounding of the reference."""
  sysname, env = VARIANTS[variant]
  g, e = util.load_golden(sysname), env_golden(variant)
  desc = util.load_desc(sysname)
  if env.kind == 'grasp':  # the affine action map derived from the descriptor == the env's own
    np.testing.assert_allclose(env.act_min, e['act_min'], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(env.act_range, e['act_range'], rtol=1e-12, atol=1e-12)
  for dtype, tol in ((np.float64, 1e-9), (np.float32, None)):
    o = Oracle(desc, dtype)
    pos_in, act_in = env_oracle.env_prologue(desc, env, g['pos'], g['act'], dtype)   # grasp.py:75-85
    out = o.step(pos_in, g['rot'], g['vel'], g['ang'], act_in)
    fa, fv = env_oracle.flat_angle_vel(desc, *o.angle_vel(out['rot'], out['ang']))
    r = env_oracle.env_epilogue(desc, env, g['pos'], out, g['act'], fa, fv, dtype)
    if tol is not None:
      for k in ('obs', 'reward', 'done', 'metrics'):
        np.testing.assert_allclose(r[k], e[k], rtol=0, atol=tol, err_msg=k)
      if env.kind in ('fetch', 'grasp'):
        # the step's output QP, hit targets teleported with the uniforms the reference drew
        out = dict(out, pos=env_oracle.env_teleport(env, out['pos'], r['metrics'][:, 0], e['rand'], dtype))
        for k in util.QP_KEYS:
          np.testing.assert_allclose(out[k], e['qp_' + k], rtol=0, atol=1e-9, err_msg=k)
        if variant == 'fetch':
          assert (e['metrics'][:, 0] > 0).any()   # the golden does hold hits
    else:
      np.testing.assert_array_equal(r['done'], e['done'])
      # velocities are position differences over dt (ant.py:226): ulp(pos)/dt of float32
      np.testing.assert_allclose(r['reward'], e['reward'], rtol=1e-4, atol=5e-3)
      assert np.quantile(np.abs(r['obs'] - e['obs']), 0.99) < 1e-3


def test_obs_sizes():
  assert VARIANTS['ant'][1].obs_size(util.load_desc('ant')) == 87
  assert VARIANTS['ant_nocfrc'][1].obs_size(util.load_desc('ant')) == 27
  assert VARIANTS['humanoid'][1].obs_size(util.load_desc('humanoid')) == 240
  assert VARIANTS['humanoid_spring'][1].obs_size(util.load_desc('humanoid_spring')) == 227
  assert VARIANTS['fetch'][1].obs_size(util.load_desc('fetch')) == 101
  assert VARIANTS['grasp'][1].obs_size(util.load_desc('grasp')) == 132


@pytest.mark.parametrize('fname', ['ant_wrapped_env.npz', 'ant_wrapped_repeat_env.npz'])
def test_wrapper_oracle_matches_reference(fname):
  """EpisodeWrapper + AutoResetWrapper (wrappers.py:85-148), the reference's own classes run
  around its Ant (oracle/gen_golden_wrappers.py): the oracle's wrapped step reproduces every
  stored step -- done by the healthy range, truncation at episode_length, the step counter
  reset after a done, state and obs replaced by first_qp / first_obs."""
  g = np.load(os.path.join(util.GOLDEN, fname))
  desc, env = util.load_desc('ant'), VARIANTS['ant'][1]
  o = Oracle(desc, np.float64)
  assert g['done'].sum() >= 3 and g['truncation'].sum() >= 2 and (g['done'] > g['truncation']).any()
  nseq, t_n = g['done'].shape
  for t in range(t_n):
    qp = {k: g[k][:, t] for k in util.QP_KEYS}
    first = {k: g['first_' + k] for k in util.QP_KEYS}
    r = env_oracle.wrapped_step(desc, env, o, qp, g['done_in'][:, t], g['steps_in'][:, t], g['act'][:, t], first,
                                g['first_obs'], int(g['episode_length']), np.float64, action_repeat=int(g['action_repeat']))
    for k in util.QP_KEYS:
      np.testing.assert_allclose(r['qp'][k], g[k + '_o'][:, t], rtol=0, atol=1e-9, err_msg='%s step %d' % (k, t))
    for k in ('obs', 'reward', 'done', 'steps', 'truncation', 'metrics'):
      np.testing.assert_allclose(r[k], g[k][:, t], rtol=0, atol=1e-9, err_msg='%s step %d' % (k, t))


def test_reset_obs_oracle_matches_reference():
  """The obs `Env.reset` returns (`_get_obs(qp, sys.info(qp))`, tests/golden/reset_obs.npz from
  the reference's own env classes): the oracle's `info_contact` reproduces the reference's
  `sys.info(qp).contact` (system.py:327-340 / :378-390) on every draw -- zero for most, non-zero for two
  humanoid draws and one fetch draw where a collider sits in the plane moving downwards -- and the
  oracle's reset obs built on it equals the reference's obs on EVERY draw."""
  g = np.load(os.path.join(util.GOLDEN, 'reset.npz'))
  go = np.load(os.path.join(util.GOLDEN, 'reset_obs.npz'))
  nonzero = 0
  for name in ('ant', 'ant_nocfrc', 'ant_spring', 'humanoid', 'humanoid_spring', 'fetch', 'grasp'):
    sysname, env = VARIANTS[name]
    desc = util.load_desc(sysname)
    qp = {k: g[sysname + '/' + k] for k in util.QP_KEYS}
    o = Oracle(desc, np.float64)
    ref = go[name + '/obs']
    ic = o.info_contact(qp['pos'], qp['rot'], qp['vel'], qp['ang'])
    np.testing.assert_allclose(ic['contact_vel'], go[name + '/contact_vel'], rtol=0, atol=1e-12, err_msg=name)
    np.testing.assert_allclose(ic['contact_ang'], go[name + '/contact_ang'], rtol=0, atol=1e-12, err_msg=name)
    nonzero += int((np.abs(go[name + '/contact_vel']).max(axis=(1, 2)) > 0).sum())
    full = env_oracle.reset_obs(desc, env, o, qp, np.float64)
    np.testing.assert_allclose(full, ref, rtol=0, atol=1e-12, err_msg=name)
  assert nonzero >= 3


def test_eval_wrapper_oracle_matches_reference():
  """EvalWrapper (wrappers.py:168-200, the reference's class around the wrapped Ant of
  gen_golden_wrappers.py): accumulated episode metrics, active flags and episode steps."""
  g = np.load(os.path.join(util.GOLDEN, 'ant_wrapped_env.npz'))
  nseq, t_n = g['done'].shape
  m = g['metrics'].shape[-1]
  em, act, es = np.zeros((nseq, m + 1)), np.ones(nseq), np.zeros(nseq)   # reset: wrappers.py:173-180
  for t in range(t_n):
    em, act, es = env_oracle.eval_step(em, act, es, g['metrics'][:, t], g['reward'][:, t], g['done'][:, t], g['steps'][:, t])
    np.testing.assert_allclose(em, g['eval_episode_metrics'][:, t], rtol=0, atol=1e-12)
    np.testing.assert_array_equal(act, g['eval_active_episodes'][:, t])
    np.testing.assert_array_equal(es, g['eval_episode_steps'][:, t])
  assert g['eval_active_episodes'][:, -1].sum() == 0 and g['eval_active_episodes'][:, 0].sum() == 2
