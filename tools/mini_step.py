"""Tiny driver for compute-sanitizer: a few steps of one system on 64 envs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, load_system  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 64
s = load_system(name)
if len(sys.argv) > 3:
  s.configure(int(sys.argv[3]))
q = s.default_qp(batch=b, device='cuda')
rng = np.random.default_rng(0)
for t in range(3):
  act = torch.from_numpy(rng.uniform(-1, 1, (b, s.act_size)).astype(np.float32)).cuda()
  q, info = s.step(q, act)
torch.cuda.synchronize()
print(name, 'ok', float(q.pos.abs().max()))
