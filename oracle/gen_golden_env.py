"""Adds tests/golden/<name>_env.npz: the REFERENCE's `Env.step` (obs, reward,
done, metrics) on the steps already stored in tests/golden/<sys>_steps.npz
(build container only; the unmodified /root/reference/brax/envs/{ant,humanoid}.py
on numpy float64 through oracle/refsys.py).

`brax.envs.__init__` imports gym/dm_env (absent here), so `brax.envs` is entered
as a bare namespace package and only the env module itself is executed:
brax/envs/ant.py:222-281 (Ant.step, _get_obs), brax/envs/humanoid.py:246-340.

Variants (name -> system golden file, env class, constructor kwargs):
  ant               ant             Ant(use_contact_forces=True)  the fork's registration
                                                                   (envs/__init__.py:45)
  ant_nocfrc        ant             Ant()                          upstream default
  ant_spring        ant_spring      Ant(use_contact_forces=True, legacy_spring=True)
  humanoid          humanoid        Humanoid()
  humanoid_spring   humanoid_spring Humanoid(legacy_spring=True)

Stored per variant: obs [S, O], reward [S], done [S], metrics [S, M] in the order
of `metric_names`, and the physics outputs of the same call (checked against the
stored system golden so both files describe the same steps).

usage: python oracle/gen_golden_env.py [name ...]
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402

VARIANTS = {
    'ant': ('ant', 'ant', 'Ant', dict(use_contact_forces=True)),
    'ant_nocfrc': ('ant', 'ant', 'Ant', dict()),
    'ant_spring': ('ant_spring', 'ant', 'Ant', dict(use_contact_forces=True, legacy_spring=True)),
    'humanoid': ('humanoid', 'humanoid', 'Humanoid', dict()),
    'humanoid_spring': ('humanoid_spring', 'humanoid', 'Humanoid', dict(legacy_spring=True)),
    # egocentric envs; their random target teleport on a hit (JAX PRNG) is the caller's.  Grasp
    # maps the action and moves the palm BEFORE sys.step, so its physics is not that of the
    # stored system golden: the step's output QP is stored here as well.
    'fetch': ('fetch', 'fetch', 'Fetch', dict()),
    'grasp': ('grasp', 'grasp', 'Grasp', dict()),
}


def env_module(name):
  brax = refsys.setup()
  if 'brax.envs' not in sys.modules:
    pkg = types.ModuleType('brax.envs')
    pkg.__path__ = [os.path.join(refsys.REFERENCE, 'brax', 'envs')]
    sys.modules['brax.envs'] = pkg
    brax.envs = pkg
  import importlib
  return importlib.import_module('brax.envs.' + name)


def main(names):
  out_dir = os.path.join(ROOT, 'tests', 'golden')
  for name in names:
    sysname, modname, clsname, kw = VARIANTS[name]
    mod = env_module(modname)
    import brax
    from brax.envs import env as benv
    e = getattr(mod, clsname)(**kw)
    g = np.load(os.path.join(out_dir, sysname + '_steps.npz'))
    steps = g['pos'].shape[0]
    obs, rew, done, met, names_m = [], [], [], [], None
    qpo = {k: [] for k in ('pos', 'rot', 'vel', 'ang')}
    rands = []
    for t in range(steps):
      qp = brax.QP(pos=g['pos'][t], rot=g['rot'][t], vel=g['vel'][t], ang=g['ang'][t])
      st = benv.State(qp=qp, obs=np.zeros(1), reward=np.zeros(()), done=np.zeros(()), metrics={},
                      info={'rng': np.array([1, 2], dtype='uint32')})
      st2 = e.step(st, g['act'][t])
      if modname != 'grasp':  # same physics as the stored system golden
        np.testing.assert_allclose(np.asarray(st2.qp.pos, np.float64)[:12], g['pos_o'][t][:12], rtol=0, atol=1e-12)
      for k in qpo:
        qpo[k].append(np.asarray(getattr(st2.qp, k), np.float64))
      if modname in ('fetch', 'grasp'):
        # the uniforms `_random_target` drew from state.info['rng'] (fetch.py:124-134,
        # grasp.py:188-198): same jumpy calls on the same key
        import brax.jumpy as jp
        keys = jp.random_split(np.array([1, 2], dtype='uint32'), 3 if modname == 'fetch' else 4)
        u = [float(jp.random_uniform(k)) for k in keys[1:]] + [0.0]
        rands.append(u[:3])
      obs.append(np.asarray(st2.obs, np.float64))
      rew.append(float(st2.reward))
      done.append(float(st2.done))
      if names_m is None:
        names_m = list(st2.metrics.keys())
      met.append([float(st2.metrics[k]) for k in names_m])
    np.savez_compressed(os.path.join(out_dir, name + '_env.npz'), obs=np.stack(obs),
                        reward=np.asarray(rew), done=np.asarray(done), metrics=np.asarray(met),
                        metric_names=np.asarray(names_m), system=np.asarray(sysname),
                        dt=np.asarray(float(e.sys.config.dt)),
                        **{'qp_' + k: np.stack(v) for k, v in qpo.items()},
                        **({'rand': np.asarray(rands)} if rands else {}),
                        **({'act_min': np.asarray(e._min_act, np.float64), 'act_range': np.asarray(e._range_act, np.float64)}
                           if modname == 'grasp' else {}))
    print('%-16s %3d steps, obs %d, metrics %s' % (name, steps, obs[0].shape[0], names_m))


if __name__ == '__main__':
  main(sys.argv[1:] or list(VARIANTS))
