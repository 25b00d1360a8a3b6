"""Kernel-only timing for A/B work on the step kernel (GPU box).

usage: [BRAX_B200_LIB=path.so] python tools/kbench.py [env:B[:threads[:contact_section]] ...] [--steps K] [--check]

Per case: K chained `System.step` launches behind a spin kernel, ONE event pair around them
(bench.py's method), rotating over enough env batches to exceed L2; prints us/step and M env-steps/s.
--check also prints max |kernel - float32 oracle| of one step on 256 harvested envs (test
infrastructure: the oracle is the checker here, as in tests/).
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, load_system  # noqa: E402


def state(s, b, seed):
  rng = np.random.default_rng(seed)
  q = s.default_qp(batch=b)
  mp = np.asarray([x['pos_mask'] for x in s.desc.bodies], np.float32)
  mr = np.asarray([x['rot_mask'] for x in s.desc.bodies], np.float32)
  vel = rng.uniform(-.1, .1, q.pos.shape).astype(np.float32) * mp
  ang = rng.uniform(-.1, .1, q.pos.shape).astype(np.float32) * mr
  return QP(*[torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (q.pos, q.rot, vel, ang)])


def run(name, b, threads, steps, check, cs=0):
  s = load_system(name)
  if threads or cs:
    s.configure(threads, cs)
  nsets = max(2, -(-2 * 126_000_000 // (b * s.desc.bytes_per_env_step())))
  nsets = min(nsets, 64)
  qps = [state(s, b, 1 + i) for i in range(nsets)]
  acts = torch.rand((steps + 20, b, s.act_size), device='cuda') * 2 - 1
  for i in range(20):
    j = i % nsets
    qps[j], _ = s.step(qps[j], acts[i])
  torch.cuda.synchronize()
  best = 1e9
  for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(int((20e-3 + 0.25e-3 * steps) * 1.9e9))
    e0.record()
    for i in range(steps):
      j = i % nsets
      qps[j], _ = s.step(qps[j], acts[20 + i])
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / steps)
  fin = bool(torch.isfinite(qps[0].pos).all())
  line = '%-16s B=%-7d T=%-3d cs=%d %8.2f us/step %8.2f M env-steps/s finite=%s' % (
      name, b, int(s._lib.bx_system_threads(s._h)), cs, best * 1e3, b / best / 1e3, fin)
  if check:
    from oracle.oracle import Oracle
    q = qps[0]
    n = min(b, 256)
    ins = [getattr(q, k)[:n].contiguous() for k in ('pos', 'rot', 'vel', 'ang')]
    a = acts[0, :n].contiguous()
    out, _ = s.step(QP(*ins), a)
    hin = [x.cpu().numpy() for x in ins]
    o32 = Oracle(s.desc, np.float32).step(*hin, a.cpu().numpy())
    o64 = Oracle(s.desc, np.float64).step(*hin, a.cpu().numpy())
    errs = []
    for k in ('pos', 'rot', 'vel', 'ang'):
      g = getattr(out, k).cpu().numpy()
      e = np.abs(g - o32[k]).reshape(n, -1).max(1)
      gap = np.abs(o32[k] - o64[k]).reshape(n, -1).max(1)
      e64 = np.abs(g - o64[k]).reshape(n, -1).max(1)
      errs.append('%s k-o32 med %.1e max %.1e | k-o64 med %.1e | o32-o64 med %.1e max %.1e' % (
          k, np.median(e), e.max(), np.median(e64), np.median(gap), gap.max()))
    line += '\n    ' + '\n    '.join(errs)
  print(line, flush=True)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('cases', nargs='*', default=['ant:8192', 'humanoid:8192', 'grasp:4096', 'fetch:8192', 'ant:1024'])
  ap.add_argument('--steps', type=int, default=100)
  ap.add_argument('--check', action='store_true')
  args = ap.parse_args()
  print('lib:', os.environ.get('BRAX_B200_LIB', 'default'))
  for c in args.cases:
    p = c.split(':')
    run(p[0], int(p[1]), int(p[2]) if len(p) > 2 else 0, args.steps, args.check, int(p[3]) if len(p) > 3 else 0)


if __name__ == '__main__':
  main()
