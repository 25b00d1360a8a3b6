"""Per-field, per-env error / tolerance of one golden system under the parity policy (GPU box).
usage: python tools/parity_case.py NAME [...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, System  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from tests import util  # noqa: E402
from tests.test_gpu_parity import ACCUMULATORS, GAP_MULT, VEL_ULPS, _env_gap, _flat  # noqa: E402

for name in sys.argv[1:]:
  g = util.load_golden(name)
  d = util.load_desc(name)
  s = System(d)
  ins = [g[k].astype(np.float32) for k in util.QP_KEYS]
  act = g['act'].astype(np.float32)
  qo, info = s.step(QP(*ins), act)
  got = _flat(qo, info)
  ref = Oracle(d, np.float32).step(*ins, act)
  r64 = Oracle(d, np.float64).step(*ins, act)
  b = ref['pos'].shape[0]
  floor = VEL_ULPS * 6e-8 / d.dt_sub
  print('==', name, 'B', b, 'dt_sub', d.dt_sub)
  for k in util.QP_KEYS + util.INFO_KEYS:
    if ref[k].size == 0:
      continue
    err = np.abs(got[k] - ref[k])
    mag = np.abs(ref[k])
    if k in ACCUMULATORS:
      mag = np.broadcast_to(mag.reshape(b, -1).max(axis=1).reshape((b,) + (1,) * (mag.ndim - 1)), mag.shape)
    tol = 1e-6 + 1e-5 * mag + GAP_MULT * _env_gap(ref[k], r64[k])
    if k not in ('pos', 'rot', 'contact_pos', 'contact_normal', 'contact_penetration'):
      tol = tol + floor
    ratio = (err / tol).reshape(b, -1).max(axis=1)
    e64 = np.abs(got[k] - r64[k]).reshape(b, -1).max(axis=1)
    gap = _env_gap(ref[k], r64[k]).reshape(b)
    print('  %-20s max err/tol per env %s | err %s | k-o64 %s | o32-o64 %s' % (
        k, np.round(ratio, 2), np.array2string(err.reshape(b, -1).max(axis=1), precision=2),
        np.array2string(e64, precision=2), np.array2string(gap, precision=2)))
