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
