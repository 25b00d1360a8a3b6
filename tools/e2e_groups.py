"""e2e of the host path with G independent env groups in flight (GPU box).
usage: [BX_HOST_STAGGER_GBPS=..] python tools/e2e_groups.py ENV B G [steps]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, load_system  # noqa: E402

name, b, G = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 50
s = load_system(name)
rng = np.random.default_rng(0)
q0 = s.default_qp(batch=b)
vel = rng.uniform(-.1, .1, q0.pos.shape).astype(np.float32)
acts = rng.uniform(-1, 1, (8, b, s.act_size)).astype(np.float32)
edges = [b * g // G for g in range(G + 1)]
chains = []
for g in range(G):
  lo, hi = edges[g], edges[g + 1]
  hb = [s.host_buffers(hi - lo), s.host_buffers(hi - lo)]
  hb[0].out['pos_o'][...] = q0.pos[lo:hi]; hb[0].out['rot_o'][...] = q0.rot[lo:hi]
  hb[0].out['vel_o'][...] = vel[lo:hi]; hb[0].out['ang_o'][...] = vel[lo:hi]
  chains.append({'hb': hb, 'lo': lo, 'hi': hi,
                 'qp': QP(hb[0].out['pos_o'], hb[0].out['rot_o'], hb[0].out['vel_o'], hb[0].out['ang_o'])})


def issue(g, i):
  c = chains[g]
  c['hb'][i % 2].act[...] = acts[i % 8, c['lo']:c['hi']]
  c['pending'] = s.step_async(c['qp'], c['hb'][i % 2].act, out=c['hb'][(i + 1) % 2], slot=g)


def finish(g):
  s.wait(g)
  chains[g]['qp'] = chains[g]['pending'][0]


for rep in range(2):
  t0 = time.perf_counter()
  for g in range(G):
    issue(g, 0)
  for i in range(1, steps):
    for g in range(G):
      finish(g)
      issue(g, i)
  for g in range(G):
    finish(g)
  el = time.perf_counter() - t0
h2d, d2h = s.host_transfer_bytes(b)
print('%s B=%d G=%d stagger=%s: %.1f M env-steps/s  %.3f ms/step  d2h %.1f GB/s h2d %.1f GB/s finite=%s' % (
    name, b, G, os.environ.get('BX_HOST_STAGGER_GBPS', 'default'), b * steps / el / 1e6, 1e3 * el / steps,
    d2h * steps / el / 1e9, h2d * steps / el / 1e9, np.isfinite(chains[0]['qp'].pos).all()))
