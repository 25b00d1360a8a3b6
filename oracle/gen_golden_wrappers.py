"""Adds tests/golden/ant_wrapped_env.npz: the REFERENCE's EpisodeWrapper + AutoResetWrapper
(brax/envs/wrappers.py:85-148) around its Ant env, as `envs.create('ant', episode_length=...,
auto_reset=True)` composes them (brax/envs/__init__.py:74-90), on numpy float64 (build
container only).  wrappers.py imports gym / dm_env at module level, so the two class
definitions are taken out of the file's AST and compiled unmodified against the reference's
own `brax.envs.env` and `brax.jumpy`.

Stored: SEQ sequences of T steps.  Each starts from the reference's own reset (first_qp /
first_obs) with state.qp replaced by a stored golden state; sequence 1 lifts the torso out of
the healthy range so `done` fires and the state is reset; episode_length = 4 makes the
truncation path fire at step 4.  Per step: the state before (qp, done, steps), the action and
the state after (qp, obs, reward, done, steps, truncation, metrics).

usage: python oracle/gen_golden_wrappers.py
"""
import ast
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402
from gen_golden_env import env_module  # noqa: E402

EPISODE_LENGTH = 4
T = 6


def wrapper_classes():
  env_module('ant')   # sets up the bare brax.envs package
  from brax import jumpy as jp
  from brax.envs import env as brax_env
  path = os.path.join(refsys.REFERENCE, 'brax', 'envs', 'wrappers.py')
  src = open(path).read()
  tree = ast.parse(src)
  keep = [n for n in tree.body if isinstance(n, ast.ClassDef)
          and n.name in ('EpisodeWrapper', 'AutoResetWrapper', 'EvalMetrics', 'EvalWrapper')]
  mod = ast.Module(body=keep, type_ignores=[])
  import typing
  import flax
  import jax
  ns = {'brax_env': brax_env, 'jp': jp, 'copy': copy, 'flax': flax, 'jax': jax, 'Dict': typing.Dict}
  exec(compile(mod, path, 'exec'), ns)
  return ns['EpisodeWrapper'], ns['AutoResetWrapper'], ns['EvalWrapper']


def main():
  generate('ant_wrapped_env.npz', 1)
  # EpisodeWrapper's action_repeat (wrappers.py:106-111): two env steps per wrapped step
  generate('ant_wrapped_repeat_env.npz', 2)


def generate(fname, action_repeat):
  out_dir = os.path.join(ROOT, 'tests', 'golden')
  Episode, AutoReset, Eval = wrapper_classes()
  import brax
  import brax.jumpy as jp
  ant = env_module('ant').Ant(use_contact_forces=True)
  # eval_metrics=True wraps EvalWrapper outermost (brax/envs/__init__.py:88-89); it only adds
  # state.info['eval_metrics'] and metrics['reward'], so the sequences are those of the inner stack
  env = Eval(AutoReset(Episode(ant, EPISODE_LENGTH, action_repeat)))
  g = np.load(os.path.join(out_dir, 'ant_steps.npz'))
  rng = np.random.default_rng(5)
  names = None
  rec = {k: [] for k in ('pos', 'rot', 'vel', 'ang', 'done_in', 'steps_in', 'act', 'pos_o', 'rot_o', 'vel_o',
                         'ang_o', 'obs', 'reward', 'done', 'steps', 'truncation', 'metrics',
                         'eval_episode_metrics', 'eval_active_episodes', 'eval_episode_steps')}
  firsts = {k: [] for k in ('first_pos', 'first_rot', 'first_vel', 'first_ang', 'first_obs')}
  for seq in range(3):
    st = env.reset(jp.random_prngkey(seq))
    for k, a in zip(('first_pos', 'first_rot', 'first_vel', 'first_ang'), ('pos', 'rot', 'vel', 'ang')):
      firsts[k].append(np.asarray(getattr(st.info['first_qp'], a), np.float64))
    firsts['first_obs'].append(np.asarray(st.info['first_obs'], np.float64))
    i = 5 * seq
    pos = np.array(g['pos'][i])
    if seq == 1:
      pos[:, 2] += 0.6      # torso above healthy_z_range -> done at the first step
    st = st.replace(qp=brax.QP(pos=pos, rot=g['rot'][i], vel=g['vel'][i], ang=g['ang'][i]))
    for t in range(T):
      act = rng.uniform(-1, 1, 8)
      for k in ('pos', 'rot', 'vel', 'ang'):
        rec[k].append(np.asarray(getattr(st.qp, k), np.float64))
      rec['done_in'].append(float(st.done))
      rec['steps_in'].append(float(st.info['steps']))
      rec['act'].append(act)
      st = env.step(st, act)
      for k in ('pos', 'rot', 'vel', 'ang'):
        rec[k + '_o'].append(np.asarray(getattr(st.qp, k), np.float64))
      rec['obs'].append(np.asarray(st.obs, np.float64))
      rec['reward'].append(float(st.reward))
      rec['done'].append(float(st.done))
      rec['steps'].append(float(st.info['steps']))
      rec['truncation'].append(float(st.info['truncation']))
      names = names or [k for k in st.metrics.keys() if k != 'reward']
      rec['metrics'].append([float(st.metrics[k]) for k in names])
      em = st.info['eval_metrics']                                   # wrappers.py:186-200
      rec['eval_episode_metrics'].append([float(em.episode_metrics[k]) for k in names + ['reward']])
      rec['eval_active_episodes'].append(float(em.active_episodes))
      rec['eval_episode_steps'].append(float(em.episode_steps))
  shape = lambda v: np.asarray(v).reshape((3, T) + np.asarray(v[0]).shape)  # noqa: E731
  np.savez_compressed(os.path.join(out_dir, fname),
                      episode_length=np.asarray(EPISODE_LENGTH), action_repeat=np.asarray(action_repeat),
                      metric_names=np.asarray(names),
                      **{k: shape(v) for k, v in rec.items()}, **{k: np.stack(v) for k, v in firsts.items()})
  d = np.asarray(rec['done']).reshape(3, T)
  print('done\n', d, '\ntruncation\n', np.asarray(rec['truncation']).reshape(3, T), '\nsteps\n',
        np.asarray(rec['steps']).reshape(3, T))


if __name__ == '__main__':
  main()
