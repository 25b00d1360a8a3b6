"""The acting loop of the reference's PPO (`brax/training/acting.py:53-74`) on the CUDA path, three ways.

    python examples/acting_loop.py [env] [num_envs] [unroll]

1. the reference's shape: `action = policy(obs); state = env.step(state, action)` -- a torch policy between single-step
   launches (every `Env.step` of ant / humanoid / fetch / grasp is ONE launch: physics + obs / reward / done / metrics,
   with EpisodeWrapper + AutoResetWrapper folded in);
2. the same policy evaluated INSIDE the launch (`bx_policy`): `unroll` steps per launch, nothing but the final state, the
   per-step obs / reward / done and the chosen actions ever leave the kernel;
3. (2) under domain randomisation: per-env collider friction and joint sockets
   (`experimental/tracing/randomizers.py`: friction_randomizer, joint_randomizer).

Needs a CUDA device (there is no CPU fallback).  Weights are random: this shows the plumbing and prints throughput.
"""
import sys
import time

import torch

sys.path.insert(0, __file__.rsplit('/', 2)[0])
from brax_b200 import envs  # noqa: E402
from brax_b200.system import AutoReset, Policy  # noqa: E402


def main():
  name = sys.argv[1] if len(sys.argv) > 1 else 'ant'
  b = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
  unroll = int(sys.argv[3]) if len(sys.argv) > 3 else 10
  dev = torch.device('cuda', 0)
  env = envs.create(name, episode_length=1000, auto_reset=True)          # envs.create(name, episode_length, auto_reset)
  no, na = env.observation_size, env.action_size
  sizes = (no, 32, 32, 32, 32, 2 * na)                                   # ppo/networks.py:67-69
  params = [((torch.rand(sizes[i], sizes[i + 1], device=dev) * 2 - 1) * (3.0 / sizes[i]) ** 0.5,
             torch.zeros(sizes[i + 1], device=dev)) for i in range(len(sizes) - 1)]

  def policy_torch(obs, noise):                                          # networks.MLP + NormalTanhDistribution.sample
    x = obs
    for i, (w, bias) in enumerate(params):
      x = x @ w + bias
      if i + 1 < len(params):
        x = x * torch.sigmoid(x)
    return torch.tanh(x[:, :na] + (torch.nn.functional.softplus(x[:, na:]) + 0.001) * noise)

  # 1. policy in torch, one launch per env step
  state = env.reset(b, device=dev, seed=0)
  noise = torch.randn(unroll, b, na, device=dev)
  for rep in range(4):                                                   # (the first pass warms cuBLAS and the allocator)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for t in range(unroll):
      state = env.step(state, policy_torch(state.obs, noise[t]))
    torch.cuda.synchronize()
    t1 = time.perf_counter() - t0
  print('%s x %d: torch policy + one launch per step   %7.1f M env-steps/s' % (name, b, b * unroll / t1 / 1e6))

  # 2. the policy inside the launch, `unroll` steps per launch
  pol = Policy(params, activation='swish')
  sysm, inner = env.env.sys, env.env
  ar = AutoReset(state.info['first_qp'], state.info['first_obs'], state.done, state.info['steps'], 1000)
  kw = dict(env=True, policy=pol, obs0=state.obs, noise=noise, steps=unroll, auto_reset=ar, info=False)
  if hasattr(inner, '_uniform'):                                        # fetch / grasp: uniforms of the target teleport
    kw['env_rand'] = torch.rand(unroll, b, 3, device=dev)
  reps = 8
  for rep in range(2):                                                   # (the first pass warms torch's allocator)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
      qp, _, out = sysm.rollout(state.qp, None, **kw)
    torch.cuda.synchronize()
    t2 = (time.perf_counter() - t0) / reps
  print('%s x %d: policy inside the launch, T = %-3d     %7.1f M env-steps/s   (reward mean %.3f, %d episodes ended)' % (
      name, b, unroll, b * unroll / t2 / 1e6, float(out['reward'].mean()), int(out['done'].sum())))

  # 3. ... with per-env friction and joint sockets
  friction = torch.empty(b, device=dev).uniform_(0.4, 1.6)
  sockets = torch.from_numpy(sysm.joint_offsets()).to(dev)[None] + torch.empty(b, 1, 3, device=dev).uniform_(-0.01, 0.01)
  qp, _, out = sysm.rollout(state.qp, None, friction=friction, joint_offset=sockets.contiguous(), **kw)
  torch.cuda.synchronize()
  print('%s x %d: ... under friction / joint-socket randomisation: reward mean %.3f, finite %s' % (
      name, b, float(out['reward'].mean()), bool(torch.isfinite(qp.pos).all())))


if __name__ == '__main__':
  main()
