"""Cycles per phase of CTA 0 (GPU box).  Builds a -DBX_PHASE_TIMING copy of the
library into build/ and runs a few steps.  usage: phase_timing.py ENV [B] [T] [contact_section]"""
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
if os.environ.get('BX_PT_LIB'):   # a prebuilt timing library (experiments)
  out = os.path.join(ROOT, os.environ['BX_PT_LIB'])
else:
  bb.build(extra=['-DBX_PHASE_TIMING'], out=out, force='--build-only' in sys.argv)
if '--build-only' in sys.argv:
  sys.exit(0)
bl.LIB_PATH = out
from brax_b200.system import load_system  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s = load_system(name)
if len(sys.argv) > 3:
  s.configure(int(sys.argv[3]), int(sys.argv[4]) if len(sys.argv) > 4 else 0)
lib = bl.load()
dbg = torch.zeros(32 + 4 * 4096 + 64, dtype=torch.int64, device='cuda')
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
names = ['tables', 'load', 'T', 'B1', 'J', 'B2', 'C1|fused', 'B3', 'C2', 'B4', 'tail', 'store', 'stage', 'flush']
tot = v[:14].sum()
print('%s B=%d: %.0f cycles/step (%.1f us @1.965GHz)' % (name, b, tot, tot / 1965))
for i, (nm, x) in enumerate(zip(names, v)):
  print('  %-9s %8.0f  %5.1f%%   warp 0 busy %8.0f' % (nm, x, 100 * x / tot, v[16 + i]))

mk = v[32 + 4 * 4096:]
if mk.any():
  print('within-item markers (warp 0, cycles/step):', [int(x) for x in mk[:16]], [int(x) for x in mk[16:32]])
# per-CTA records of the LAST launch: tiles per SM, duration by co-residency, start skew
ntile = min((b + 31) // 32, 4096)
rec = dbg.cpu().numpy()[32:32 + 4 * ntile].reshape(ntile, 4)
sm, t0, t1, cyc = rec[:, 0], rec[:, 1], rec[:, 2], rec[:, 3]
start = t0.min()
per_sm = np.bincount(sm.astype(np.int64))
print('tiles %d on %d SMs; tiles/SM histogram %s' % (ntile, (per_sm > 0).sum(), np.bincount(per_sm[per_sm > 0]).tolist()))
print('kernel span (first CTA start -> last CTA end): %.1f us' % ((t1.max() - start) / 1e3))
print('CTA start skew: median %.1f us, max %.1f us' % (np.median(t0 - start) / 1e3, (t0.max() - start) / 1e3))
for n_co in sorted(set(per_sm[per_sm > 0].tolist())):
  m = per_sm[sm.astype(np.int64)] == n_co
  print('  %d tile(s)/SM: %4d CTAs, duration median %.1f us (min %.1f max %.1f), %.0f cycles' % (
      n_co, m.sum(), np.median((t1 - t0)[m]) / 1e3, (t1 - t0)[m].min() / 1e3, (t1 - t0)[m].max() / 1e3, np.median(cyc[m])))
print('CTA 0 on SM %d with %d tile(s)' % (sm[0], per_sm[int(sm[0])]))
# event time of single launches (enqueued behind a spin kernel) vs the CTA span above
evs = []
for a in acts[:10]:
  torch.cuda._sleep(int(300e-6 * 1.9e9))
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  q, _ = s.step(q, a)
  e1.record()
  evs.append((e0, e1))
torch.cuda.synchronize()
rec = dbg.cpu().numpy()[32:32 + 4 * ntile].reshape(ntile, 4)
print('event-timed launch: median %.1f us; CTA span of the last one %.1f us' % (
    1e3 * np.median([x.elapsed_time(y) for x, y in evs]), (rec[:, 2].max() - rec[:, 1].min()) / 1e3))
# two launches back to back inside one event pair: (pair - single) = steady-state cost
torch.cuda._sleep(int(300e-6 * 1.9e9))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for a in acts[:8]:
  q, _ = s.step(q, a)
e1.record()
torch.cuda.synchronize()
print('8 launches in one event pair: %.1f us per launch' % (1e3 * e0.elapsed_time(e1) / 8))
