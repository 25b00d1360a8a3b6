"""Compact accuracy summary over all golden systems (GPU box): per system and QP field,
max |kernel - o64| next to max |o32 - o64| (the float32 oracle's own error).
usage: [BRAX_B200_LIB=...] python tools/err_summary.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, System  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from tests import util  # noqa: E402

print('lib', os.environ.get('BRAX_B200_LIB', 'default'))
tot = {k: [] for k in util.QP_KEYS}
for name in util.golden_names():
  g = util.load_golden(name)
  d = util.load_desc(name)
  s = System(d)
  ins = [g[k].astype(np.float32) for k in util.QP_KEYS]
  act = g['act'].astype(np.float32)
  qo, _ = s.step(QP(*ins), act)
  o32 = Oracle(d, np.float32).step(*ins, act)
  o64 = Oracle(d, np.float64).step(*ins, act)
  parts = []
  for k in util.QP_KEYS:
    e = np.abs(np.asarray(getattr(qo, k)) - o64[k]).max()
    r = np.abs(o32[k] - o64[k]).max()
    tot[k].append(e / max(r, 1e-12))
    parts.append('%s %.1e/%.1e' % (k, e, r))
  print('%-26s %s' % (name, '  '.join(parts)))
print('geomean ratio kernel/oracle32 error:', {k: round(float(np.exp(np.mean(np.log(v)))), 2) for k, v in tot.items()})
