"""TEST INFRASTRUCTURE ONLY: exact operation count of one `System.step`.

Builds oracle/brax_step_oracle.c as C++ with the counting scalar of
oracle/flopcount.hpp and steps a few representative environments (states
rolled forward from a noisy reset so contacts are active).  Writes
oracle/flops.json: per system, operations per env-step
  flops = add + mul + div + sqrt + trig (each counted 1; an FMA would be 2),
with the categories listed so other weightings can be derived.  bench.py's
FP32 roofline uses `flops`.

usage: python oracle/count_flops.py
"""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from brax_b200.desc import SystemDesc  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

NAMES = ['add', 'mul', 'div', 'sqrt', 'trig', 'cmp', 'minmax']


def build():
  lib = os.path.join(HERE, 'liboracle_count.so')
  src = os.path.join(HERE, 'brax_step_oracle.c')
  if not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(src):
    subprocess.check_call(['g++', '-x', 'c++', '-O1', '-fPIC', '-shared', '-DBXO_COUNT',
                           '-ffp-contract=off', '-Wno-unused-function', src, '-o', lib, '-lm'])
  return lib


def count(name, envs=64, warm=20):
  d = SystemDesc.load(os.path.join(ROOT, 'brax_b200', 'systems', name + '.json'))
  o = Oracle(d, np.float32)
  rng = np.random.default_rng(0)
  q = d.default_qp
  pos = np.broadcast_to(np.asarray(q['pos'], np.float32), (envs, d.num_bodies, 3)).copy()
  rot = np.broadcast_to(np.asarray(q['rot'], np.float32), (envs, d.num_bodies, 4)).copy()
  mp = np.asarray([x['pos_mask'] for x in d.bodies], np.float32)
  mr = np.asarray([x['rot_mask'] for x in d.bodies], np.float32)
  st = {'pos': pos, 'rot': rot,
        'vel': rng.uniform(-.1, .1, pos.shape).astype(np.float32) * mp,
        'ang': rng.uniform(-.1, .1, pos.shape).astype(np.float32) * mr}
  for _ in range(warm):
    act = rng.uniform(-1, 1, (envs, d.act_size)).astype(np.float32)
    st = o.step(st['pos'], st['rot'], st['vel'], st['ang'], act)
  # same struct layout (CR wraps one float): reuse the float32 oracle's packing
  lib = C.CDLL(build())
  counters = (C.c_double * len(NAMES)).in_dll(lib, 'bxo_counters')
  for i in range(len(NAMES)):
    counters[i] = 0.0
  act = rng.uniform(-1, 1, (envs, d.act_size)).astype(np.float32)
  o._step = lambda s, b, io, nt: lib.bxo_step_count(s, b, io)
  o.step(st['pos'], st['rot'], st['vel'], st['ang'], act)
  c = {n: counters[i] / envs for i, n in enumerate(NAMES)}
  c['flops'] = c['add'] + c['mul'] + c['div'] + c['sqrt'] + c['trig']
  c['envs_sampled'] = envs
  return c


def main():
  out = {}
  for name in ('ant', 'humanoid', 'grasp', 'fetch', 'ant_spring', 'humanoid_spring'):
    out[name] = count(name)
    print(name, {k: round(v, 1) for k, v in out[name].items()})
  with open(os.path.join(HERE, 'flops.json'), 'w') as f:
    json.dump(out, f, indent=1)


if __name__ == '__main__':
  main()
