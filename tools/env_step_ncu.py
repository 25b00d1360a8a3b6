"""Driver for an ncu capture of the env build of the step kernel: Env.reset, then wrapped env steps
(episode + auto-reset bookkeeping in the launch) at 8192 envs.  usage: env_step_ncu.py [env] [B]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200 import envs  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
env = envs.create(name, episode_length=1000, auto_reset=True)
st = env.reset(b, device='cuda', seed=0)
st.info['steps'] = torch.randint(0, 1000, (b,), device='cuda').float()
for t in range(8):
  st = env.step(st, torch.rand(b, env.action_size, device='cuda') * 2 - 1)
torch.cuda.synchronize()
print(name, 'ok', float(st.reward.mean()))
