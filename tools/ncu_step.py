"""Short driver for ncu captures of the step kernel: 40 chained `System.step` launches of one env at
the bench workload, rotating over env batches larger than L2 (as bench.py does).
usage: python tools/ncu_step.py [env] [B]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200.system import QP, load_system  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s = load_system(name)
rng = np.random.default_rng(0)
nsets = max(2, -(-2 * 126_000_000 // (b * s.desc.bytes_per_env_step())))
q0 = s.default_qp(batch=b)
mp = np.asarray([x['pos_mask'] for x in s.desc.bodies], np.float32)
qps = []
for i in range(nsets):
  vel = rng.uniform(-.1, .1, q0.pos.shape).astype(np.float32) * mp
  qps.append(QP(*[torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (q0.pos, q0.rot, vel, vel)]))
acts = torch.rand((40, b, s.act_size), device='cuda') * 2 - 1
for t in range(40):
  j = t % nsets
  qps[j], _ = s.step(qps[j], acts[t])
torch.cuda.synchronize()
print(name, b, 'ok', bool(torch.isfinite(qps[0].pos).all()))
