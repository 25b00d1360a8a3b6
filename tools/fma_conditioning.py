"""How well-conditioned is a system's step in float32?  (CPU only)

Builds the C oracle twice in float32 -- as the tests use it (-ffp-contract=off) and
with multiply-adds fused (-ffp-contract=fast -mfma) -- and counts the environments
in which these two *equally valid* float32 evaluations of the same step differ by
more than the parity policy of tests/test_gpu_parity.py allows.  That count is the
floor of the kernel's own outlier count (the kernel fuses multiply-adds too).
usage: python tools/fma_conditioning.py [name:steps ...]
"""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from tests import util  # noqa: E402

ACC = ('contact_vel', 'contact_ang', 'joint_vel', 'joint_ang', 'actuator_vel', 'actuator_ang')


def harvest(d, b, steps, seed):
  rng = np.random.default_rng(seed)
  q = d.default_qp
  pos = np.broadcast_to(np.asarray(q['pos'], np.float32), (b, d.num_bodies, 3)).copy()
  rot = np.broadcast_to(np.asarray(q['rot'], np.float32), (b, d.num_bodies, 4)).copy()
  mp = np.asarray([x['pos_mask'] for x in d.bodies], np.float32)
  mr = np.asarray([x['rot_mask'] for x in d.bodies], np.float32)
  vel = rng.uniform(-.3, .3, pos.shape).astype(np.float32) * mp
  ang = rng.uniform(-.3, .3, pos.shape).astype(np.float32) * mr
  o = O.Oracle(d, np.float32)
  for _ in range(steps):
    r = o.step(pos, rot, vel, ang, rng.uniform(-1, 1, (b, d.act_size)).astype(np.float32))
    pos, rot, vel, ang = r['pos'], r['rot'], r['vel'], r['ang']
  return (pos, rot, vel, ang), rng


def main():
  td = tempfile.mkdtemp()
  lib = os.path.join(td, 'liboracle_f32_fma.so')
  subprocess.check_call(['gcc', '-O2', '-fPIC', '-shared', '-fopenmp', '-ffp-contract=fast', '-mfma', '-mavx2',
                         '-Wno-unused-function', O.SRC, '-o', lib, '-lm'])
  cases = [a.split(':') for a in sys.argv[1:]] or [('ant', 30), ('humanoid', 30), ('grasp', 20), ('fetch', 20),
                                                   ('half_cheetah', 6), ('humanoid_standup', 20)]
  b = 1024
  for name, steps in cases:
    d = util.load_desc(name)
    ins, rng = harvest(d, b, int(steps), 5)
    act = rng.uniform(-1, 1, (b, d.act_size)).astype(np.float32)
    o32, o64, ofm = O.Oracle(d, np.float32), O.Oracle(d, np.float64), O.Oracle(d, np.float32)
    ofm._lib = C.CDLL(lib)
    ofm._step = getattr(ofm._lib, o32._step.__name__)
    ofm._step.restype = C.c_int
    ref, r64, got = o32.step(*ins, act), o64.step(*ins, act), ofm.step(*ins, act)
    floor = 4 * 6e-8 / d.dt_sub
    bad = np.zeros(b, bool)
    for k in util.QP_KEYS + util.INFO_KEYS:
      if ref[k].size == 0:
        continue
      err, mag = np.abs(got[k] - ref[k]), np.abs(ref[k])
      if k in ACC:
        mag = np.broadcast_to(mag.reshape(b, -1).max(axis=1).reshape((b,) + (1,) * (mag.ndim - 1)), mag.shape)
      g = np.abs(ref[k].astype(np.float64) - r64[k])
      g = g.reshape(b, -1).max(axis=1).reshape((-1,) + (1,) * (g.ndim - 1))
      tol = 1e-6 + 1e-5 * mag + 4 * g
      if k not in ('pos', 'rot', 'contact_pos', 'contact_normal', 'contact_penetration'):
        tol = tol + floor
      bad |= (err > tol).reshape(b, -1).any(axis=1)
    print('%-18s %2d steps in: fused vs unfused float32 outside the policy in %3d of %d envs (%.1f%%), max|ang in| %.0f rad/s'
          % (name, int(steps), bad.sum(), b, 100.0 * bad.sum() / b, np.abs(ins[3]).max()))


if __name__ == '__main__':
  main()
