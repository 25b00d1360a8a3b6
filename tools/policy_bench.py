"""Timing of the in-kernel policy rollout (GPU box): env rollout with actions from memory vs with bx_policy.

usage: [BRAX_B200_LIB=path.so] python tools/policy_bench.py [env:B ...] [--t 10] [--launches 40]
Prints the median launch time per step of (a) `rollout(env=True)` fed actions[T,B,A], (b) the same launch with
the PPO default policy network inside (deterministic), (c) with sampled actions (noise[T,B,A]).
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from brax_b200 import envs as bx_envs  # noqa: E402
from brax_b200.system import Policy  # noqa: E402


def median_ms(fn, n):
  for i in range(3):
    fn(i)
  torch.cuda.synchronize()
  ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
  for i, (a, b) in enumerate(ev):
    a.record()
    fn(i)
    b.record()
  torch.cuda.synchronize()
  return sorted(a.elapsed_time(b) for a, b in ev)[n // 2]


def run(name, b, t, n, hidden):
  env = bx_envs.create(name)
  s = env.sys
  na, no = env.action_size, env.observation_size
  nsets = 8
  qps = [env.reset(b, device='cuda', seed=i).qp for i in range(nsets)]
  obs = [s.observe(q) for q in qps]
  acts = torch.rand(t, b, na, device='cuda') * 2 - 1
  sizes = (no,) + hidden + (2 * na,)
  pol = Policy([((torch.rand(sizes[i], sizes[i + 1], device='cuda') * 2 - 1) * (3.0 / sizes[i]) ** 0.5,
                 torch.zeros(sizes[i + 1], device='cuda')) for i in range(len(sizes) - 1)], activation='swish',
               obs_mean=torch.zeros(no, device='cuda'), obs_std=torch.ones(no, device='cuda'))
  noise = torch.randn(t, b, na, device='cuda')

  def plain(i):
    j = i % nsets
    qps[j], _, eo = s.rollout(qps[j], acts, env=True, info=False)

  def with_policy(nz):
    def f(i):
      j = i % nsets
      qps[j], _, eo = s.rollout(qps[j], None, env=True, policy=pol, obs0=obs[j], noise=nz, steps=t, info=False)
      obs[j] = eo['obs'][-1]
    return f

  a = median_ms(plain, n) / t
  d = median_ms(with_policy(None), n) / t
  z = median_ms(with_policy(noise), n) / t
  print('%-10s B=%-6d T=%d hidden=%s  actions from memory %7.2f us/step | policy (mode) %7.2f (+%.2f) | policy (sample) '
        '%7.2f (+%.2f)  -> %.1f M env-steps/s' % (name, b, t, 'x'.join(map(str, hidden)), a * 1e3, d * 1e3, (d - a) * 1e3,
                                                  z * 1e3, (z - a) * 1e3, b / z / 1e3), flush=True)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('cases', nargs='*', default=['ant:8192', 'humanoid:8192', 'ant:4096'])
  ap.add_argument('--t', type=int, default=10)
  ap.add_argument('--launches', type=int, default=40)
  ap.add_argument('--hidden', default='32,32,32,32')
  args = ap.parse_args()
  print('lib:', os.environ.get('BRAX_B200_LIB', 'default'))
  for c in args.cases:
    name, b = c.split(':')
    run(name, int(b), args.t, args.launches, tuple(int(x) for x in args.hidden.split(',')))


if __name__ == '__main__':
  main()
