"""Tiny driver (ncu launch lists): Env.reset (bx_default_qp + observe), env steps (fused epilogue),
a short rollout and a wrapped (episode + auto-reset) rollout on 70 envs (two full tiles + a ragged one).  usage: mini_env_step.py ant|humanoid|fetch|grasp [legacy_spring]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200 import envs  # noqa: E402
from brax_b200.envs import State  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
spring = len(sys.argv) > 2
b = 70
env = envs.create(name, legacy_spring=spring)
st = env.reset(batch=b, device='cuda', velocity_noise=0.1)
rng = np.random.default_rng(0)
for t in range(3):
  act = torch.from_numpy(rng.uniform(-1, 1, (b, env.action_size)).astype(np.float32)).cuda()
  st = env.step(st, act)
acts = torch.from_numpy(rng.uniform(-1, 1, (3, b, env.action_size)).astype(np.float32)).cuda()
kw = {}
if name in ('fetch', 'grasp'):
  kw['env_rand'] = torch.rand((3, b, 3), device='cuda')
qp, info, eo = env.sys.rollout(st.qp, acts, env=True, **kw)
w = envs.EpisodeAutoReset(env, 2)
last, ew = w.rollout(w.reset(b, device='cuda', seed=1), acts)
torch.cuda.synchronize()
assert float(ew['truncation'].sum()) > 0
print(name, 'ok', float(eo['obs'].abs().max()), float(eo['reward'].mean()))
