"""Dumps the envs where kernel and float32 oracle disagree most (GPU box).
usage: python tools/outliers.py NAME B STEPS OUT.npz
Saves inputs and the kernel / o32 / o64 outputs of the 8 worst envs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402
from tests.test_gpu_parity import _flat, _harvest, _oracle, _system  # noqa: E402

name, b, steps, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
s = _system(name)
qp, rng = _harvest(s, b, steps, 5)
act = rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32)
got = _flat(*s.step(qp, torch.from_numpy(act).cuda()))
ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
o32 = _oracle(name, np.float32).step(*ins, act)
o64 = _oracle(name, np.float64).step(*ins, act)
err = np.abs(got['rot'] - o32['rot']).reshape(b, -1).max(1) + np.abs(got['pos'] - o32['pos']).reshape(b, -1).max(1)
idx = np.argsort(-err)[:8]
print('worst envs', idx, err[idx])
for i in idx:
  print(i, 'k-o32 %.2e k-o64 %.2e o32-o64 %.2e' % (
      np.abs(got['rot'][i] - o32['rot'][i]).max(), np.abs(got['rot'][i] - o64['rot'][i]).max(),
      np.abs(o32['rot'][i] - o64['rot'][i]).max()),
      'finite in', np.isfinite(ins[0][i]).all(), 'max|vel|', np.abs(ins[2][i]).max(), 'max|ang|', np.abs(ins[3][i]).max())
sv = {'idx': idx, 'act': act[idx]}
for k, v in zip(util.QP_KEYS, ins):
  sv['in_' + k] = v[idx]
for k in got:
  sv['k_' + k] = got[k][idx]; sv['o32_' + k] = o32[k][idx]; sv['o64_' + k] = o64[k][idx]
np.savez(out, **sv)
