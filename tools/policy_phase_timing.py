"""Cycles inside the in-kernel policy (CTA 0, warp 0; GPU box): builds the -DBX_PHASE_TIMING library
(tools/phase_timing.py) and runs policy rollouts.  usage: policy_phase_timing.py [ENV] [B] [hidden,...]"""
import ctypes as C
import os
import sys

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
from brax_b200 import envs as bx_envs  # noqa: E402
from brax_b200.system import Policy  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
hidden = tuple(int(x) for x in (sys.argv[3] if len(sys.argv) > 3 else '32,32,32,32').split(','))
env = bx_envs.create(name)
s = env.sys
lib = bl.load()
na, no = env.action_size, env.observation_size
sizes = (no,) + hidden + (2 * na,)
pol = Policy([((torch.rand(sizes[i], sizes[i + 1], device='cuda') * 2 - 1) * (3.0 / sizes[i]) ** 0.5,
               torch.zeros(sizes[i + 1], device='cuda')) for i in range(len(sizes) - 1)], activation='swish',
             obs_mean=torch.zeros(no, device='cuda'), obs_std=torch.ones(no, device='cuda'))
t = 10
noise = torch.randn(t, b, na, device='cuda')
qp = env.reset(b, device='cuda', seed=0).qp
obs = s.observe(qp)
for i in range(3):
  qp, _, eo = s.rollout(qp, None, env=True, policy=pol, obs0=obs, noise=noise, steps=t, info=False)
  obs = eo['obs'][-1].clone()
torch.cuda.synchronize()
dbg = torch.zeros(32 + 4 * 4096 + 64, dtype=torch.int64, device='cuda')
lib.bx_debug_set_buffer.argtypes = [C.c_void_p]
lib.bx_debug_set_buffer(C.c_void_p(dbg.data_ptr()))
n = 5
for i in range(n):
  qp, _, eo = s.rollout(qp, None, env=True, policy=pol, obs0=obs, noise=noise, steps=t, info=False)
  obs = eo['obs'][-1].clone()
torch.cuda.synchronize()
v = dbg.cpu().numpy() / (n * t)
mk = v[32 + 4 * 4096:]
print('%s B=%d hidden=%s: cycles per step of warp 0 / CTA 0 inside policy_forward' % (name, b, hidden))
print('  obs pass + barrier %6.0f' % mk[16])
for l in range(len(sizes) - 1):
  print('  layer %d (%3d -> %3d): compute %6.0f  barrier %6.0f' % (l, sizes[l], sizes[l + 1], mk[32 + l], mk[17 + l]))
print('  distribution + barrier %6.0f' % mk[23])
print('  total %6.0f cycles = %.2f us @1.965 GHz' % (mk[16:48].sum(), mk[16:48].sum() / 1965))
