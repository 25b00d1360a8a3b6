"""Times the reset path on the GPU: bx_default_qp (forward kinematics) and the observe launch, and the
whole Env.reset, with CUDA events.  usage: reset_timing.py [envs per launch]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200 import envs  # noqa: E402

b = int(sys.argv[1]) if len(sys.argv) > 1 else 8192


def timed(fn, n=50):
  for _ in range(5):
    fn()
  torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(n):
    fn()
  e1.record()
  torch.cuda.synchronize()
  return e0.elapsed_time(e1) / n * 1e3


for name in ('ant', 'humanoid', 'fetch', 'grasp'):
  env = envs.create(name)
  s = env.sys
  nd = s.num_joint_dof
  ja = torch.from_numpy(s.default_angle()).cuda() + (torch.rand(b, nd, device='cuda') * 2 - 1) * 0.1
  jv = (torch.rand(b, nd, device='cuda') * 2 - 1) * 0.1
  qp = s.default_qp(joint_angle=ja, joint_velocity=jv)
  t_fk = timed(lambda: s.default_qp(joint_angle=ja, joint_velocity=jv))
  t_obs = timed(lambda: s.observe(qp))
  t_all = timed(lambda: env.reset(b, device='cuda', seed=1))
  byts = 4 * (2 * nd + 13 * s.num_bodies)
  print('%-9s B=%d  default_qp %.1f us (%.0f GB/s of %d B/env)  observe %.1f us  Env.reset %.1f us' % (
      name, b, t_fk, b * byts / t_fk / 1e3, byts, t_obs, t_all))
