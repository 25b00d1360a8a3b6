"""The committed descriptors (brax_b200/systems/*.json) ARE what `brax_b200.extract.describe`
reads off the reference's constructed `brax.System` -- the drop-in route `System.from_brax`
(INTEGRATION.md 2).  Needs the reference tree (build container only; skipped elsewhere)."""
import json
import os
import sys

import pytest

from tests import util

sys.path.insert(0, os.path.join(util.ROOT, 'oracle'))
import refsys  # noqa: E402

pytestmark = pytest.mark.skipif(not os.path.isdir(refsys.REFERENCE), reason='reference tree not present')


@pytest.mark.parametrize('name,env,var,extra', [
    ('ant', 'ant', '_SYSTEM_CONFIG', 0), ('humanoid_spring', 'humanoid', '_SYSTEM_CONFIG_SPRING', 0),
    ('grasp', 'grasp', '_SYSTEM_CONFIG', 3), ('fetch', 'fetch', '_SYSTEM_CONFIG', 0)])
def test_committed_descriptor_is_a_fresh_extraction(name, env, var, extra):
  from brax_b200.extract import describe
  s = refsys.env_system(env, var)
  d = describe(s, name, act_size=int(s.num_joint_dof + s.num_forces_dof) + extra)
  d.validate()
  want = json.load(open(os.path.join(util.SYSTEMS, name + '.json')))
  assert json.loads(d.to_json()) == want
  # the reset tables describe the same default state the reference returns
  assert d.reset['num_dof'] == len(s.default_angle())


class _OracleBackend:
  """`brax_b200.System.step`'s signature on numpy, computed by the float64 oracle -- stands in for the CUDA
  backend so that the shim's plumbing can run where there is no GPU (test infrastructure only)."""

  def __init__(self, desc):
    import numpy as np
    try:
      from oracle.oracle import Oracle
    except ImportError:      # oracle/ itself is on sys.path here (refsys): `oracle` is then oracle/oracle.py
      from oracle import Oracle
    self._o, self.act_size, self._np = Oracle(desc, np.float64), desc.act_size, np

  def step(self, qp, act):
    from brax_b200.system import Info, P, QP
    o = self._o.step(qp.pos[None], qp.rot[None], qp.vel[None], qp.ang[None], self._np.asarray(act)[None])
    o = {k: v[0] for k, v in o.items()}
    return (QP(o['pos'], o['rot'], o['vel'], o['ang']),
            Info(P(o['contact_vel'], o['contact_ang']), P(o['joint_vel'], o['joint_ang']),
                 P(o['actuator_vel'], o['actuator_ang']), o['contact_pos'], o['contact_normal'], o['contact_penetration']))


@pytest.mark.parametrize('variant', ['ant', 'humanoid', 'grasp'])
def test_shim_installed_into_the_reference_env(variant):
  """`brax_b200.shim.install(env.sys)` overrides `step` ONLY: the reference's own `Env.step` (obs, reward, done,
  metrics built by the reference's code from the QP / Info the shim returns) gives what the unpatched reference
  gives.  Backend here = the float64 oracle; on a GPU box it is the CUDA path (no reference tree there)."""
  import numpy as np
  path0 = list(sys.path)
  try:
    import gen_golden_env as gg      # (oracle/ is on sys.path: refsys above)
  finally:
    sys.path[:] = path0              # gen_golden_env puts oracle/ in front of the repo root: later tests that spawn
                                     # processes (tests/test_shard_gloo.py) must keep resolving `oracle` as the package
  import brax_b200.shim as shim
  from brax_b200.extract import describe
  sysname, modname, clsname, kw = gg.VARIANTS[variant]
  mod = gg.env_module(modname)
  import brax
  from brax.envs import env as benv
  e = getattr(mod, clsname)(**kw)
  g = np.load(os.path.join(util.GOLDEN, sysname + '_steps.npz'))
  mk = lambda t: benv.State(qp=brax.QP(pos=g['pos'][t], rot=g['rot'][t], vel=g['vel'][t], ang=g['ang'][t]),  # noqa: E731
                            obs=np.zeros(1), reward=np.zeros(()), done=np.zeros(()), metrics={},
                            info={'rng': np.array([1, 2], dtype='uint32')})
  act_n = e.action_size
  want = [e.step(mk(t), g['act'][t][:act_n]) for t in range(2)]
  desc = describe(e.sys, sysname, act_size=util.load_desc(sysname).act_size)
  shim.install(e.sys, backend=_OracleBackend(desc))
  assert e.sys.step.__name__ == 'step' and e.sys.step._b200 is not None
  for t in range(2):
    got = e.step(mk(t), g['act'][t][:act_n])
    assert type(got.qp) is brax.QP
    # (the shim hands float32 arrays in: rounding the INPUT state alone moves a humanoid step by up to 5e-6 in
    # exact arithmetic)
    np.testing.assert_allclose(np.asarray(got.obs), np.asarray(want[t].obs), rtol=1e-5, atol=2e-5)
    np.testing.assert_allclose(float(got.reward), float(want[t].reward), rtol=1e-5, atol=2e-4)
    assert float(got.done) == float(want[t].done)
    for k in want[t].metrics:
      np.testing.assert_allclose(float(got.metrics[k]), float(want[t].metrics[k]), rtol=1e-5, atol=2e-4)
    np.testing.assert_allclose(np.asarray(got.qp.pos), np.asarray(want[t].qp.pos), rtol=0, atol=2e-6)
