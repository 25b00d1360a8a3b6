"""Pins the C oracle against the reference itself.

tests/golden/<system>_steps.npz were produced by oracle/gen_golden.py running
the UNMODIFIED reference `brax.System.step` (brax/physics/system.py:244) on its
numpy float64 backend.  The float64 build of the oracle must reproduce every
output of every stored step to rounding error; the float32 build (the checker
for the CUDA path) must agree to float32 accuracy.
"""
import os

import numpy as np
import pytest

from oracle.oracle import Oracle
from tests import util


@pytest.mark.parametrize('name', util.golden_names())
def test_oracle_f64_matches_reference(name):
  g = util.load_golden(name)
  o = Oracle(util.load_desc(name), np.float64)
  out = o.step(g['pos'], g['rot'], g['vel'], g['ang'], g['act'])
  for k in util.QP_KEYS:
    ref = g[k + '_o']
    err = np.abs(out[k] - ref).max()
    assert err <= 1e-11 * max(1.0, np.abs(ref).max()), (name, k, err)
  for k in util.INFO_KEYS:
    ref = g[k]
    if ref.size == 0:
      continue
    err = np.abs(out[k] - ref).max()
    assert err <= 1e-10 * max(1.0, np.abs(ref).max()), (name, k, err)


@pytest.mark.parametrize('name', util.golden_names())
def test_oracle_f32_close_to_reference(name):
  """float32 oracle vs float64 reference: the f32-vs-f64 gap, reported in
  SURVEY.md section 7: ~1e-6 on pos/rot, ~ulp(pos)/dt_sub on vel/ang."""
  g = util.load_golden(name)
  d = util.load_desc(name)
  o = Oracle(d, np.float32)
  out = o.step(g['pos'], g['rot'], g['vel'], g['ang'], g['act'])
  for k in ('pos', 'rot'):
    ref = g[k + '_o']
    assert np.all(np.abs(out[k] - ref) <= 1e-4 + 1e-5 * np.abs(ref)), (name, k)
  # contact flags: penetration sign must agree except within 1e-5 of zero
  ref = g['contact_penetration']
  if ref.size:
    bad = ((out['contact_penetration'] > 0) != (ref > 0)) & (np.abs(ref) > 1e-5)
    assert not bad.any(), name


def test_batch_equals_loop():
  g = util.load_golden('ant')
  o = Oracle(util.load_desc('ant'), np.float32)
  full = o.step(g['pos'], g['rot'], g['vel'], g['ang'], g['act'])
  one = o.step(g['pos'][3], g['rot'][3], g['vel'][3], g['ang'][3], g['act'][3])
  for k in util.QP_KEYS + util.INFO_KEYS:
    np.testing.assert_array_equal(full[k][3], one[k][0])


def _angle_vel_golden(name):
  return np.load(os.path.join(util.GOLDEN, name + '_angle_vel.npz'))


def select_free(x, dof, free):
  """The free_dofs selection of Joint.angle_vel (joints.py:219-225): of joint i keep
  its first free[i] dofs, groups in order.  x: [..., nj, 3] -> [..., sum(free)]."""
  cols = [x[..., i, :free[i]] for i in range(len(dof))]
  return np.concatenate(cols, axis=-1)


def angle_tol(angle64, dof):
  """float32 tolerance of joint angles: 2e-5, except the second angle of a 3-dof
  joint, which is acos(cos theta) (joints.py:405-407): d(theta) = d(cos)/sin(theta),
  so ~1e-6 of rounding in the cosine becomes 1e-6/|sin theta| near theta = 0."""
  tol = np.full(angle64.shape, 2e-5)
  three = np.asarray(dof) == 3
  tol[..., three, 1] += 1e-6 / np.maximum(np.abs(np.sin(angle64[..., three, 1])), 1e-3)
  return tol


@pytest.mark.parametrize('name', util.golden_names())
def test_oracle_angle_vel_matches_reference(name):
  """bxo_angle_vel (float64) reproduces the reference's Joint.angle_vel on every
  stored output state, per joint and after the free_dofs selection the envs see
  (tests/golden/*_angle_vel.npz, oracle/gen_golden_angle_vel.py)."""
  g, av = util.load_golden(name), _angle_vel_golden(name)
  o = Oracle(util.load_desc(name), np.float64)
  angle, vel = o.angle_vel(g['rot_o'], g['ang_o'])
  np.testing.assert_allclose(angle, av['angle'], rtol=0, atol=1e-11)
  np.testing.assert_allclose(vel, av['vel'], rtol=0, atol=1e-11)
  np.testing.assert_allclose(select_free(angle, av['dof'], av['free']), av['flat_angle'], rtol=0, atol=1e-11)
  np.testing.assert_allclose(select_free(vel, av['dof'], av['free']), av['flat_vel'], rtol=0, atol=1e-11)
  # the float32 build (the CUDA path's checker) stays within float32 rounding of it
  a32, v32 = Oracle(util.load_desc(name), np.float32).angle_vel(g['rot_o'], g['ang_o'])
  assert (np.abs(a32 - av['angle']) <= angle_tol(av['angle'], av['dof'])).all()
  np.testing.assert_allclose(v32, av['vel'], rtol=1e-5, atol=2e-5)
