"""Cycles per phase of CTA 0 (GPU box).  Builds a -DBX_PHASE_TIMING copy of the
library into build/ and runs a few steps.  usage: phase_timing.py ENV [B] [T]"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200 import build as bb, lib as bl  # noqa: E402

out = os.path.join(ROOT, 'build', 'libbrax_b200_pt.so')
os.makedirs(os.path.dirname(out), exist_ok=True)
bb.build(extra=['-DBX_PHASE_TIMING'], out=out, force='--build-only' in sys.argv)
if '--build-only' in sys.argv:
  sys.exit(0)
bl.LIB_PATH = out
from brax_b200.system import load_system  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s = load_system(name)
if len(sys.argv) > 3:
  s.configure(int(sys.argv[3]))
lib = bl.load()
dbg = torch.zeros(16, dtype=torch.int64, device='cuda')
lib.bx_debug_set_buffer.argtypes = [C.c_void_p]
q = s.default_qp(batch=b, device='cuda')
rng = np.random.default_rng(0)
acts = [torch.from_numpy(rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32)).cuda() for _ in range(30)]
for a in acts[:20]:
  q, _ = s.step(q, a)
torch.cuda.synchronize()
lib.bx_debug_set_buffer(C.c_void_p(dbg.data_ptr()))
n = 10
for a in acts[20:]:
  q, _ = s.step(q, a)
torch.cuda.synchronize()
v = dbg.cpu().numpy() / n
names = ['tables', 'load', 'T', 'B1', 'J', 'B2', 'C1', 'B3', 'C2', 'B4', 'tail', 'store', 'stage', 'flush']
tot = v[:14].sum()
print('%s B=%d: %.0f cycles/step (%.1f us @1.965GHz)' % (name, b, tot, tot / 1965))
for nm, x in zip(names, v):
  print('  %-7s %8.0f  %5.1f%%' % (nm, x, 100 * x / tot))
