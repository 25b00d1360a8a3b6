"""Pins the C oracle against the reference itself.

tests/golden/<system>_steps.npz were produced by oracle/gen_golden.py running
the UNMODIFIED reference `brax.System.step` (brax/physics/system.py:244) on its
numpy float64 backend.  The float64 build of the oracle must reproduce every
output of every stored step to rounding error; the float32 build (the checker
for the CUDA path) must agree to float32 accuracy.
"""
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
