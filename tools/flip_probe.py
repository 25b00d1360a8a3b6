"""Finds the env where kernel and oracles disagree most in the full-size test scenario and shows
where along the substeps the two part (GPU box).  usage: python tools/flip_probe.py NAME B"""
import copy
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, System  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402
from tests import util  # noqa: E402
from tests.test_gpu_parity import _flat, _harvest, _system  # noqa: E402

name, b = sys.argv[1], int(sys.argv[2])
s = _system(name)
d = s.desc
qp, rng = _harvest(s, b, 20, 21)
act = torch.from_numpy(rng.uniform(-1, 1, (b, d.act_size)).astype(np.float32)).cuda()
got = _flat(*s.step(qp, act))
ins = [getattr(qp, k).cpu().numpy() for k in util.QP_KEYS]
a = act.cpu().numpy()
o32 = Oracle(d, np.float32).step(*ins, a)
o64 = Oracle(d, np.float64).step(*ins, a)
err = np.abs(got['rot'] - o32['rot']).reshape(b, -1).max(1)
gap = np.abs(o32['rot'] - o64['rot']).reshape(b, -1).max(1)
score = err / (1e-6 + 4 * gap)
worst = np.argsort(-score)[:3]
print('worst envs', worst, 'rot err', err[worst], 'o32-o64', gap[worst])
for w in worst[:2]:
  one = [x[w:w + 1] for x in ins]
  print('env', w)
  for n in range(2, d.substeps + 1, 2):
    d2 = copy.deepcopy(d)
    d2.substeps = n
    s2 = System(d2)
    k = _flat(*s2.step(QP(*[torch.from_numpy(x).cuda() for x in one]), act[w:w + 1]))
    r32 = Oracle(d2, np.float32).step(*one, a[w:w + 1])
    r64 = Oracle(d2, np.float64).step(*one, a[w:w + 1])
    print('  substeps %d: rot k-o32 %.2e o32-o64 %.2e | ang k-o32 %.2e | actuator_ang k-o32 %.2e (max %.1e) | contact_vel k-o32 %.2e | pen flags k %s o32 %s' % (
        n, np.abs(k['rot'] - r32['rot']).max(), np.abs(r32['rot'] - r64['rot']).max(), np.abs(k['ang'] - r32['ang']).max(),
        np.abs(k['actuator_ang'] - r32['actuator_ang']).max(), np.abs(r32['actuator_ang']).max(),
        np.abs(k['contact_vel'] - r32['contact_vel']).max(),
        (k['contact_penetration'][0] > 0).astype(int), (r32['contact_penetration'][0] > 0).astype(int)))
    if n == 2:
      ang, _ = Oracle(d2, np.float64).angle_vel(one[1], one[3])
      lim = np.array([j['limit'] + [0] * (6 - len(j['limit'])) for j in d.joints]).reshape(-1, 3, 2)
      print('  input joint angles (deg):', np.round(np.degrees(ang[0]), 3).tolist())
      print('  dist to limits (deg):', np.round(np.degrees(np.minimum(np.abs(ang[0] - lim[..., 0]), np.abs(ang[0] - lim[..., 1]))), 4).tolist())
  np.savez(os.path.join(ROOT, 'gpurun_out', 'flip_%s_%d.npz' % (name, w)), act=a[w], **{k: v[w] for k, v in zip(util.QP_KEYS, ins)},
           **{'k_' + k: v[w] for k, v in got.items()})
