"""The numpy restatement of `System.default_qp(joint_angle, joint_velocity)`
(oracle/reset_oracle.py, system.py:112-242) against the reference's own outputs for all 31 env
configs (tests/golden/reset.npz, oracle/gen_golden_reset.py)."""
import os

import numpy as np
import pytest

from oracle import reset_oracle
from tests import util

G = np.load(os.path.join(util.GOLDEN, 'reset.npz'))
NAMES = sorted({k.split('/')[0] for k in G.files})


def test_every_system_has_reset_vectors():
  assert NAMES == sorted(f[:-5] for f in os.listdir(util.SYSTEMS) if f.endswith('.json'))


@pytest.mark.parametrize('name', NAMES)
def test_default_qp_matches_reference(name):
  desc = util.load_desc(name)
  assert desc.reset['num_dof'] == G[name + '/angle'].shape[1]
  np.testing.assert_allclose(desc.reset['default_angle'], G[name + '/angle'][0], rtol=0, atol=1e-15)
  got = reset_oracle.default_qp(desc, G[name + '/angle'], G[name + '/velocity'], np.float64)
  for k in util.QP_KEYS:
    np.testing.assert_allclose(got[k], G[name + '/' + k], rtol=0, atol=1e-12, err_msg=k)
  # draw 0 is default_qp() itself: the descriptor's recorded default state
  for k in util.QP_KEYS:
    np.testing.assert_allclose(got[k][0], np.asarray(desc.default_qp[k]), rtol=0, atol=1e-12, err_msg=k)
  # float32 evaluation stays within float32 rounding of it
  g32 = reset_oracle.default_qp(desc, G[name + '/angle'], G[name + '/velocity'], np.float32)
  for k in util.QP_KEYS:
    np.testing.assert_allclose(g32[k], G[name + '/' + k], rtol=1e-5, atol=2e-6, err_msg=k)
