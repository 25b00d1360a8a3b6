"""bench.py -- env-steps/s of the fused Brax v1 `System.step` on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--env ant] [--envs 8192]
  python bench.py --impl reference ...     # CPU arm (oracle port, all cores)

A "step" is one `System.step` (all substeps) over the whole batch of
environments with fresh U(-1,1) actions; the output state of step t is the
input of step t+1.  Workload = BASELINE.json configs[1]: ant, 8192 envs per
GPU, default substeps.  One JSON line is printed by rank 0.

  value     env-steps/s, state resident in HBM, CUDA events around each step's
            launch (summed), max over ranks.  Inputs larger than L2: the timed
            steps rotate over enough independent env batches (>= 2 x L2 of state)
            that every step reads cold state.  `value_l2_flushed` is the same
            measurement on ONE batch with the whole L2 (kernel code included)
            evicted by a 256 MiB fill before every step.
  e2e       the same metric through the host-buffer C-ABI call (bx_step_host
            via brax_b200.System.step on numpy views of pinned memory):
            H2D of QP+act and D2H of QP+Info inside the timed region.
  roofline  FP32-pipe bound (the path is ~50-100 flop/B): flops per
            env-step (oracle/flops.json, counted from the oracle source) x
            env-steps/s vs 148 SM x 128 lanes x 2 x sm clock; the HBM fraction
            is reported beside it from the algorithmic bytes per env-step.
  cpu_baseline  the C oracle port (float32, OpenMP over envs) on the host cores,
            bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'env-steps/sec (Brax v1 System.step, all substeps)'
UNIT = 'env-steps/s'
SM_COUNT = 148
L2_BYTES = 126 * 1000 * 1000  # B200 L2 (B200_PROFILING.md)
# dram__bytes_read.sum + dram__bytes_write.sum of one bx_step_kernel launch at the
# bench workload, from the committed `ncu --set full` capture (profiles/); bytes.
# (reads only: the 11.3 MB of outputs were still in L2 when the kernel ended)
NCU_TRAFFIC = {
    # env: (bytes per launch, capture under profiles/).  Reads + writes seen at DRAM by one launch of the
    # bench workload under ncu (--set full, cache control on: the outputs of a launch stay in the 126 MB L2
    # until a later launch evicts them, so the figure is dominated by the cold input reads).
    'ant': (4585984, 'profiles/r02_ncu_full_ant_step_kernel_final.csv (ant 8192 envs, final round-2 build)'),
    'humanoid': (5748992, 'profiles/r02_ncu_full_humanoid_step_kernel.csv (humanoid 8192 envs)'),
    'grasp': (4142848, 'profiles/r02_ncu_full_grasp_step_kernel.csv (grasp 4096 envs)'),
}
FP32_LANES = 128


def load_peaks():
  p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
  if os.path.exists(p):
    with open(p) as f:
      d = json.load(f)
    return d.get('hbm_gbs', 6650.0), d.get('sm_max_mhz', 1965.0), 'measured'
  return 6650.0, 1965.0, 'fallback'


def load_flops(env):
  p = os.path.join(ROOT, 'oracle', 'flops.json')
  if os.path.exists(p):
    with open(p) as f:
      d = json.load(f)
    if env in d:
      return d[env]
  return None


class ClockSampler:
  """Samples SM clocks / throttle reasons while the timed region runs: NVML in a
  thread (every ~2 ms, no process spawn, so the host enqueue loop is not
  disturbed), plus one sample on entry and one on exit taken by the caller's
  thread.  `nvidia-smi` (one call on exit) only if NVML cannot be loaded."""

  def __init__(self, index):
    self.index, self.rows, self._stop, self._t = index, [], False, None
    self.source, self._nv, self._hnd = 'nvml', None, None

  def _init_nvml(self):
    import pynvml as nv
    nv.nvmlInit()
    try:
      import torch
      uuid = str(torch.cuda.get_device_properties(self.index).uuid)
      hnd = nv.nvmlDeviceGetHandleByUUID(('GPU-' + uuid).encode())
    except Exception:
      hnd = nv.nvmlDeviceGetHandleByIndex(self.index)
    self._mx = nv.nvmlDeviceGetMaxClockInfo(hnd, nv.NVML_CLOCK_SM)
    self._bits = [getattr(nv, 'nvmlClocksEventReasonHwSlowdown', 0x8),
                  getattr(nv, 'nvmlClocksEventReasonHwThermalSlowdown', 0x40),
                  getattr(nv, 'nvmlClocksEventReasonSwThermalSlowdown', 0x20),
                  getattr(nv, 'nvmlClocksEventReasonSwPowerCap', 0x4)]
    self._reasons = getattr(nv, 'nvmlDeviceGetCurrentClocksEventReasons', None) or \
        nv.nvmlDeviceGetCurrentClocksThrottleReasons
    self._nv, self._hnd = nv, hnd

  def _sample(self):
    sm = self._nv.nvmlDeviceGetClockInfo(self._hnd, self._nv.NVML_CLOCK_SM)
    r = self._reasons(self._hnd)
    self.rows.append([str(sm), str(self._mx)] + ['Active' if r & bit else 'Not Active' for bit in self._bits])

  def _run(self):
    while not self._stop:
      try:
        self._sample()
      except Exception:
        pass
      time.sleep(0.002)

  def _smi_once(self):
    q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')
    try:
      out = subprocess.run(
          ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q,
           '--format=csv,noheader,nounits'], capture_output=True, text=True,
          timeout=10).stdout.strip()
      if out:
        self.rows.append([x.strip() for x in out.split(',')])
    except Exception:
      pass

  def __enter__(self):
    try:
      self._init_nvml()
      self._sample()
      self._t = threading.Thread(target=self._run, daemon=True)
      self._t.start()
    except Exception:
      self._nv, self.source = None, 'nvidia-smi'
    return self

  def __exit__(self, *a):
    self._stop = True
    if self._t is not None:
      self._t.join(timeout=2)
    if self._nv is not None:
      try:
        self._sample()
      except Exception:
        pass
    if not self.rows:
      self.source = 'nvidia-smi'
      self._smi_once()

  def summary(self):
    sm, mx, reasons = [], [], set()
    names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
    for r in self.rows:
      try:
        sm.append(float(r[0])); mx.append(float(r[1]))
      except Exception:
        continue
      for n, v in zip(names, r[2:6]):
        if v.lower().startswith('active'):
          reasons.add(n)
    return {'sm_mhz': float(np.median(sm)) if sm else None,
            'sm_max_mhz': float(max(mx)) if mx else None,
            'reasons': sorted(reasons), 'samples': len(sm), 'source': self.source}


def make_state(desc, b, seed):
  """default_qp + small velocity noise; warm-up steps then diversify it."""
  rng = np.random.default_rng(seed)
  d = desc.default_qp
  pos = np.broadcast_to(np.asarray(d['pos'], np.float32), (b, desc.num_bodies, 3)).copy()
  rot = np.broadcast_to(np.asarray(d['rot'], np.float32), (b, desc.num_bodies, 4)).copy()
  vel = rng.uniform(-.1, .1, pos.shape).astype(np.float32)
  ang = rng.uniform(-.1, .1, pos.shape).astype(np.float32)
  mask_p = np.asarray([x['pos_mask'] for x in desc.bodies], np.float32)
  mask_r = np.asarray([x['rot_mask'] for x in desc.bodies], np.float32)
  return pos, rot, vel * mask_p, ang * mask_r


def workload_name(args, desc):
  """config.workload, identical on both arms (b200 / --impl reference)."""
  return '%s %d envs/GPU, default substeps (%d), dynamics %s' % (
      args.env, args.envs, desc.substeps, desc.dynamics_mode)


def cpu_baseline(desc, env, b, budget_s=12.0, threads=0):
  """Times the oracle port (float32, OpenMP over envs) on the host cores."""
  from oracle.oracle import Oracle
  o = Oracle(desc, np.float32, baseline=True)   # -O3 -march=native build of the port (timing only)
  cores = threads or (os.cpu_count() or 1)
  rng = np.random.default_rng(7)
  pos, rot, vel, ang = make_state(desc, b, 7)
  act = rng.uniform(-1, 1, (b, desc.act_size)).astype(np.float32)
  out = o.step(pos, rot, vel, ang, act, num_threads=cores)  # warm-up
  t0, n = time.perf_counter(), 0
  while True:
    out = o.step(out['pos'], out['rot'], out['vel'], out['ang'], act, num_threads=cores)
    n += 1
    el = time.perf_counter() - t0
    if el > budget_s or n >= 5000:
      break
  return {'value': b * n / el, 'unit': UNIT, 'cores': cores, 'kind': 'port',
          'build_flags': ' '.join(o.build_flags) + ' -fopenmp',
          'sample': '%s %d envs x %d steps (%.1f s), C port of the reference float32 + OpenMP over envs '
                    '(scalar per env; the reference\'s own jit+vmap needs jax: unavailable on this box)' % (env, b, n, el)}


# The other BASELINE.json configs, measured by the default run beside the headline (configs[1]):
# (env, envs per GPU, steps) -- configs[2] humanoid 8192, configs[3] grasp 4096, configs[0] ant 128 x 1000.
SECONDARY = [('humanoid', 8192, None), ('grasp', 4096, None), ('ant', 128, 1000)]
CHUNK = 128


def timed_chunks(torch, n_steps, step_i, sync):
  """n_steps calls of step_i(i), one CUDA event pair per chunk of <= CHUNK launches queued behind a spin
  kernel that outlasts the host's enqueue time (device time only, launches chained as in a stepping
  loop); returns the summed device time in ms."""
  spin = int((20e-3 + 0.25e-3 * min(n_steps, CHUNK)) * 1.9e9)
  pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(0, n_steps, CHUNK)]
  for ci, c0 in enumerate(range(0, n_steps, CHUNK)):
    torch.cuda._sleep(spin)
    pairs[ci][0].record()
    for i in range(c0, min(n_steps, c0 + CHUNK)):
      step_i(i)
    pairs[ci][1].record()
  sync()
  return float(sum(x.elapsed_time(y) for x, y in pairs))


def measure_config(torch, name, b, steps, warmup, local, rank, sync):
  """One more config with the headline's method: `System.step` (value), the full `Env.step` in one
  launch (env_step) and 10 steps per launch (rollout).  Returns per-rank milliseconds; the caller
  reduces them (max over ranks) and turns them into whole-job env-steps/s."""
  from brax_b200 import envs as bx_envs
  from brax_b200.desc import SystemDesc
  from brax_b200.envs import State
  from brax_b200.system import QP, System
  desc = SystemDesc.load(os.path.join(ROOT, 'brax_b200', 'systems', name + '.json'))
  dev = torch.device('cuda', local)
  s = System(desc, device=local)
  bytes_set = b * desc.bytes_per_env_step()
  nsets = min(64, max(2, -(-2 * L2_BYTES // bytes_set)))
  base = QP(*[torch.from_numpy(x).to(dev) for x in make_state(desc, b, 300 + rank)])
  qps = [QP(base.pos.clone(), base.rot.clone(), base.vel.clone(), base.ang.clone()) for _ in range(nsets)]
  gen = torch.Generator(device=dev)
  gen.manual_seed(20_000 + rank)
  n_act = min(steps + warmup, 256)
  acts = torch.rand((n_act, b, desc.act_size), generator=gen, device=dev) * 2 - 1
  for i in range(max(warmup, nsets)):
    qps[i % nsets], _ = s.step(qps[i % nsets], acts[i % n_act])
  sync()

  def step_i(i):
    j = i % nsets
    qps[j], _ = s.step(qps[j], acts[(warmup + i) % n_act])
  ms = timed_chunks(torch, steps, step_i, sync)
  # T = 10 steps per launch
  roll_t = 10
  n_roll = max(1, min(steps, 100) // roll_t)
  qps[0], _ = s.rollout(qps[0], acts[:roll_t])
  er = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_roll)]
  torch.cuda._sleep(int(25e-3 * 1.9e9))
  for i in range(n_roll):
    j = i % nsets
    er[i][0].record()
    qps[j], _ = s.rollout(qps[j], acts[:roll_t])
    er[i][1].record()
  sync()
  each = sorted(x.elapsed_time(y) for x, y in er)
  ms_roll = float(each[len(each) // 2]) * n_roll
  # full Env.step in one launch
  env = bx_envs.create(name, device=local, **(dict(use_contact_forces=True) if name == 'ant' else {}))
  if name in ('fetch', 'grasp'):
    far = torch.tensor([8.0, 3.0, 1.0], device=dev)
    for q in qps:
      q.pos[:, env.env.target_body] = q.pos[:, env.env.object_body if name == 'grasp' else env.env.ref_body] + far
  sts = [State(q, None, None, None) for q in qps]
  for j in range(nsets):
    sts[j] = env.step(sts[j], acts[j % n_act])
  sync()

  def env_i(i):
    j = i % nsets
    sts[j] = env.step(sts[j], acts[(warmup + i) % n_act])
  ms_env = timed_chunks(torch, steps, env_i, sync)
  finite = bool(torch.isfinite(sts[0].qp.pos).all().item())
  info = {'env': name, 'envs_per_gpu': b, 'steps': steps, 'substeps': desc.substeps, 'dynamics': desc.dynamics_mode,
          'threads_per_tile': s_threads(s), 'env_batches': nsets, 'obs_size': env.observation_size,
          'working_set_mb': nsets * bytes_set / 1e6, 'rollout_launches': n_roll, 'finite': finite,
          'l2': ('inputs larger than L2 (%d env batches in rotation)' % nsets) if nsets * bytes_set > L2_BYTES
                else 'working set %.1f MB < L2 (a batch this small cannot exceed L2 within 64 batches)' % (nsets * bytes_set / 1e6)}
  return [ms, ms_roll, ms_env], info, desc


def config_entry(info, desc, world, ms, ms_roll, ms_env, hbm_gbs, sm_max_mhz):
  b, steps = info['envs_per_gpu'], info['steps']
  value = world * b * steps / (ms * 1e-3)
  per_gpu = value / world
  flops = load_flops(info['env'])
  peak = SM_COUNT * FP32_LANES * 2 * sm_max_mhz * 1e6 / 1e12
  out = dict(info)
  out.update({
      'workload': '%s %d envs/GPU, default substeps (%d), dynamics %s' % (info['env'], b, desc.substeps, desc.dynamics_mode),
      'value': value, 'unit': UNIT, 'ms_per_step': ms / steps,
      'roofline': {'bound': 'hbm', 'achieved': per_gpu * desc.bytes_per_env_step() / 1e9, 'peak': hbm_gbs, 'unit': 'GB/s',
                   'frac': per_gpu * desc.bytes_per_env_step() / 1e9 / hbm_gbs,
                   'bytes_per_env_step': desc.bytes_per_env_step(),
                   'traffic': NCU_TRAFFIC.get(info['env'], (None, None))[0] if b == (4096 if info['env'] == 'grasp' else 8192) else None,
                   'traffic_source': NCU_TRAFFIC.get(info['env'], (None, 'no ncu --set full capture committed'))[1]},
      'roofline_fp32': ({'bound': 'fp32', 'achieved': per_gpu * flops['flops'] / 1e12, 'peak': peak, 'unit': 'TFLOP/s',
                         'frac': per_gpu * flops['flops'] / 1e12 / peak, 'flops_per_env_step': flops['flops']} if flops else None),
      'env_step': {'value': world * b * steps / (ms_env * 1e-3), 'unit': UNIT, 'ms_per_step': ms_env / steps},
      'rollout': {'steps_per_launch': 10, 'value': world * b * 10 * info['rollout_launches'] / (ms_roll * 1e-3), 'unit': UNIT,
                  'ms_per_step': ms_roll / (10 * info['rollout_launches']), 'timing': 'median launch'},
  })
  return out



def run_sweep(args, desc):
  """BASELINE config 5: `--sweep B1,B2,...` = TOTAL batch sizes split over the ranks (contiguous env
  blocks, brax_b200.shard): per point the chunked-event device time of `--steps` steps, max over ranks;
  rank 0 prints ONE JSON line whose `sweep` holds every point (env-steps/s and roofline fractions)."""
  import torch
  import torch.distributed as dist
  from brax_b200 import shard
  from brax_b200.system import QP, System
  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  torch.cuda.set_device(local)
  dev = torch.device('cuda', local)
  s = System(desc, device=local)

  def sync():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  hbm_gbs, sm_max_mhz, _ = load_peaks()
  flops = load_flops(args.env)
  peak = SM_COUNT * FP32_LANES * 2 * sm_max_mhz * 1e6 / 1e12
  points = []
  with ClockSampler(local) as clk:
    for total in [int(x) for x in args.sweep.split(',')]:
      lo, hi = shard.shard_range(total, world, rank)
      b = hi - lo
      ms = 0.0
      if b > 0:
        bytes_set = b * desc.bytes_per_env_step()
        nsets = min(32, max(2, -(-2 * L2_BYTES // bytes_set)))
        base = QP(*[torch.from_numpy(x).to(dev) for x in make_state(desc, b, 500 + rank)])
        qps = [QP(base.pos.clone(), base.rot.clone(), base.vel.clone(), base.ang.clone()) for _ in range(nsets)]
        acts = torch.rand((16, b, desc.act_size), device=dev) * 2 - 1
        for i in range(max(args.warmup, nsets)):
          qps[i % nsets], _ = s.step(qps[i % nsets], acts[i % 16])
        sync()

        def step_i(i):
          j = i % nsets
          qps[j], _ = s.step(qps[j], acts[i % 16])
        ms = timed_chunks(torch, args.steps, step_i, sync)
        del qps, base, acts
      else:
        sync()
        sync()
      (ms,) = shard.reduce_stats([ms], ['max'], device=dev)
      v = total * args.steps / (ms * 1e-3)
      points.append({'total_envs': total, 'envs_per_gpu': -(-total // world), 'value': v, 'ms_per_step': ms / args.steps,
                     'fp32_frac': (v / world * flops['flops'] / 1e12 / peak) if flops else None,
                     'hbm_frac': v / world * desc.bytes_per_env_step() / 1e9 / hbm_gbs})
  if rank == 0:
    print(json.dumps({'metric': METRIC, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
                      'scaling': 'strong per point (total batch split over the ranks)', 'dtype': 'f32', 'data': 'synthetic',
                      'config': {'workload': '%s batch sweep, total envs over %d GPU(s), default substeps' % (args.env, world)},
                      'sweep': points, 'clocks': clk.summary()}))
  if world > 1:
    dist.destroy_process_group()


def run_reference(args, desc):
  """--impl reference: the reference's CPU implementation of the path.  JAX is
  not installable here (no network), so this is the validated C port of the
  reference's arithmetic on all host cores."""
  rank = int(os.environ.get('RANK', '0'))
  if rank != 0:
    return
  from oracle.oracle import Oracle
  o = Oracle(desc, np.float32, baseline=True)   # -O3 -march=native build of the port (timing only)
  cores = os.cpu_count() or 1
  b = args.envs  # per-step bounded sample: one GPU's shard of the workload
  rng = np.random.default_rng(7)
  pos, rot, vel, ang = make_state(desc, b, 7)
  # at least 5 untimed steps whatever --warmup says: the OpenMP team, the page cache of the outputs and the host's
  # clocks are up before the timed steps (a cold 3-step run reads 0.46 M env-steps/s instead of 0.67 M: the
  # reference arm must not look slower than it is)
  warm = max(args.warmup, 5)
  acts = rng.uniform(-1, 1, (args.steps + warm, b, desc.act_size)).astype(np.float32)
  st = {'pos': pos, 'rot': rot, 'vel': vel, 'ang': ang}
  for i in range(warm):
    st = o.step(st['pos'], st['rot'], st['vel'], st['ang'], acts[i], num_threads=cores)
  t0 = time.perf_counter()
  for i in range(args.steps):
    st = o.step(st['pos'], st['rot'], st['vel'], st['ang'], acts[warm + i], num_threads=cores)
  el = time.perf_counter() - t0
  v = b * args.steps / el
  sample = '%s %d envs per step (one GPU shard), C oracle float32 + OpenMP' % (args.env, b)
  print(json.dumps({
      'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': UNIT,
      'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
      'ms_per_step': 1e3 * el / args.steps, 'higher_is_better': True,
      'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
      'config': {'workload': workload_name(args, desc)},
      'cpu_baseline': {'value': v, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample,
                       'build_flags': ' '.join(o.build_flags) + ' -fopenmp'},
      'e2e': {'value': v, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
  }))


def run_gpu(args, desc):
  import torch
  import torch.distributed as dist
  from brax_b200 import lib
  from brax_b200.system import QP, System

  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  torch.cuda.set_device(local)
  dev = torch.device('cuda', local)
  s = System(desc, device=local)
  if args.threads or args.contact_section:
    s.configure(args.threads, args.contact_section)
  s.autotune = bool(args.autotune)
  b, a, n, k = args.envs, desc.act_size, desc.num_bodies, desc.num_contacts
  pos, rot, vel, ang = make_state(desc, b, 100 + rank)
  qp = QP(*[torch.from_numpy(x).to(dev) for x in (pos, rot, vel, ang)])
  gen = torch.Generator(device=dev)
  gen.manual_seed(10_000 + rank)
  total = args.steps + args.warmup
  acts = torch.rand((total, b, a), generator=gen, device=dev) * 2 - 1
  flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device=dev)
  # timed steps are enqueued in chunks behind a spin kernel that outlasts the host's
  # enqueue time for the chunk (~0.15 ms per step), well inside the launch queue depth
  CHUNK = 128
  spin_cycles = int((20e-3 + 0.25e-3 * min(args.steps, CHUNK)) * 1.9e9)

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  # Working set larger than L2: the timed loop rotates over `nsets` independent env
  # batches, so the state a step reads was written nsets steps earlier with
  # >= 2 x L2 of other traffic in between (cold data; kernel code and the system's
  # tables stay L2-resident, as in any real stepping loop).
  bytes_set = b * desc.bytes_per_env_step()
  nsets = max(2, -(-2 * L2_BYTES // bytes_set))
  qps = [QP(qp.pos.clone(), qp.rot.clone(), qp.vel.clone(), qp.ang.clone()) for _ in range(nsets)]
  for i in range(max(args.warmup, 0)):
    qps[i % nsets], info = s.step(qps[i % nsets], acts[i])
  for j in range(nsets):            # untimed: every set leaves its start state once
    qps[j], info = s.step(qps[j], acts[j % total])
  barrier()

  def event_pairs():
    return [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            for _ in range(args.steps)]

  def timed_chunks(step_i):
    """args.steps calls of step_i(i), one event pair per chunk of <= CHUNK launches queued
    behind a spin kernel; returns the summed device time in ms."""
    pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
             for _ in range(0, args.steps, CHUNK)]
    for ci, c0 in enumerate(range(0, args.steps, CHUNK)):
      torch.cuda._sleep(spin_cycles)
      pairs[ci][0].record()
      for i in range(c0, min(args.steps, c0 + CHUNK)):
        step_i(i)
      pairs[ci][1].record()
    barrier()
    return float(sum(x.elapsed_time(y) for x, y in pairs))

  ev, ev_fl = event_pairs(), event_pairs()
  # ---- the timed region: EXACTLY args.steps steps, device time from one CUDA event pair
  # per chunk of <= CHUNK launches.  A spin kernel holds the stream while the host enqueues
  # the chunk, so the pair holds device time only (no host enqueue gaps); no event records
  # sit between the launches, so consecutive launches chain the way they do in a real
  # stepping loop (with programmatic dependent launch the prologue of step i+1 overlaps the
  # drain of step i).
  ec = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        for _ in range(0, args.steps, CHUNK)]
  with ClockSampler(local) as clk:
    barrier()
    launches0 = lib.launch_count()
    t_wall0 = time.perf_counter()
    for ci, c0 in enumerate(range(0, args.steps, CHUNK)):
      torch.cuda._sleep(spin_cycles)
      ec[ci][0].record()
      for i in range(c0, min(args.steps, c0 + CHUNK)):
        j = i % nsets
        qps[j], info = s.step(qps[j], acts[args.warmup + i])
      ec[ci][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = lib.launch_count() - launches0   # kernels launched in the timed loop
  ms_chunk = float(sum(x.elapsed_time(y) for x, y in ec))
  # secondary: the same steps with an event pair around EVERY launch (per-step min / median /
  # max; each pair costs ~2.7 us of device time and breaks the chaining of the launches)
  for i in range(args.steps):
    if i % CHUNK == 0:
      torch.cuda._sleep(spin_cycles)
    j = i % nsets
    ev[i][0].record()
    qps[j], info = s.step(qps[j], acts[args.warmup + i])
    ev[i][1].record()
  barrier()
  qp = qps[0]
  # secondary: one env batch, whole L2 evicted (code and tables included)
  # before every step by a 256 MiB fill outside the event pair
  for i in range(args.steps):
    if i % CHUNK == 0:
      torch.cuda._sleep(2 * spin_cycles)
    flush.fill_(float(i))
    ev_fl[i][0].record()
    qp, info = s.step(qp, acts[args.warmup + i])
    ev_fl[i][1].record()
  barrier()
  # back-to-back from Python (host enqueue rate included), for reference
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  q2 = qp
  e0.record()
  for i in range(args.steps):
    q2, _ = s.step(q2, acts[args.warmup + i])
  e1.record()
  torch.cuda.synchronize()
  # the same loop without per-step host work: System.make_stepper (8 in-place steps per CUDA graph replay)
  st8 = s.make_stepper(b, steps_per_replay=8, graph=True)
  for name in ('pos', 'rot', 'vel', 'ang'):
    getattr(st8.qp, name).copy_(getattr(qp, name))
  st8.acts.copy_(acts[:8])
  st8.replay()
  n_rep = max(1, args.steps // 8)
  g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize()
  g0.record()
  for i in range(n_rep):
    st8.replay()
  g1.record()
  torch.cuda.synchronize()
  ms_graph = g0.elapsed_time(g1) / (8 * n_rep)
  # "next" row N2: T steps per launch (bx_rollout, actions [T,B,A]); the
  # state stays in shared memory between the steps of a launch.  Reported beside value.
  roll_t = min(10, args.steps)
  n_roll = max(1, args.steps // roll_t)
  er = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_roll)]
  q3, _ = s.rollout(qps[0], acts[:roll_t])   # warm-up of the multi-step launch
  torch.cuda._sleep(spin_cycles)
  for i in range(n_roll):
    j = i % nsets
    a0 = args.warmup + (i * roll_t) % max(1, args.steps - roll_t + 1)
    er[i][0].record()
    qps[j], _ = s.rollout(qps[j], acts[a0:a0 + roll_t])
    er[i][1].record()
  barrier()
  # median launch x launches: one multi-millisecond hiccup of the box inside these few launches
  # (seen about once per ten runs) would otherwise dominate the figure
  ms_roll_each = sorted(x.elapsed_time(y) for x, y in er)
  ms_roll = float(ms_roll_each[len(ms_roll_each) // 2]) * n_roll
  # "next" row N1, first piece: Joint.angle_vel of the new state as a kernel epilogue
  # (what every env's _get_obs asks for after the step); same timing method as `value`.
  for j in range(nsets):   # untimed: the allocator sees the larger output block of every set once
    qps[j], _, _ = s.step(qps[j], acts[j % total], angle_vel=True)
  barrier()

  def av_step(i):
    j = i % nsets
    qps[j], _, _ = s.step(qps[j], acts[args.warmup + i], angle_vel=True)
  ms_av = timed_chunks(av_step)
  # "next" row N1: the whole Env.step (sys.step + obs / reward / done / metrics of
  # envs/ant.py:222-281, envs/humanoid.py:246-340) as ONE launch; same timing method.
  ms_env, ms_wenv, ms_reset, env_obs, ms_pol = 0.0, 0.0, 0.0, 0, 1.0
  last_reward = None
  env_name = args.env[:-len('_spring')] if args.env.endswith('_spring') else args.env
  if env_name in ('ant', 'humanoid', 'fetch', 'grasp') and args.env != 'grasp_spring':
    from brax_b200 import envs as bx_envs
    from brax_b200.envs import State
    kw = dict(use_contact_forces=True) if env_name == 'ant' else {}   # the fork's registration
    env = bx_envs.create(env_name, legacy_spring=args.env.endswith('_spring'), device=local, **kw)
    if args.threads or args.contact_section:
      env.sys.configure(args.threads, args.contact_section)
    env_obs = env.observation_size
    if env_name in ('fetch', 'grasp'):
      # targets away from the robot, as after the reference's reset (no teleports in the timed loop)
      far = torch.tensor([8.0, 3.0, 1.0], device=dev)
      for q in qps:
        q.pos[:, env.env.target_body] = q.pos[:, env.env.object_body if env_name == 'grasp' else env.env.ref_body] + far
    sts = [State(q, None, None, None) for q in qps]
    for j in range(nsets):
      sts[j] = env.step(sts[j], acts[j % total])
    barrier()
    def env_step(i):
      j = i % nsets
      sts[j] = env.step(sts[j], acts[args.warmup + i])
    ms_env = timed_chunks(env_step)
    last_reward = sts[0].reward
    # ... and with EpisodeWrapper(1000) + AutoResetWrapper applied inside the same launch
    # (wrappers.py:105-148); step counters spread over the episode as in steady-state training,
    # so ~0.1 % of the envs end their episode and reload their first state every step
    wenv = bx_envs.EpisodeAutoReset(env, 1000)
    first_obs = torch.zeros(b, env_obs, device=dev)
    for j in range(nsets):
      q = sts[j].qp
      first = type(q)(*[getattr(q, k).clone() for k in ('pos', 'rot', 'vel', 'ang')])
      sts[j] = State(q, first_obs, None, torch.zeros(b, device=dev), {},
                     dict(first_qp=first, first_obs=first_obs,
                          steps=torch.randint(0, 1000, (b,), device=dev).float()))
      sts[j] = wenv.step(sts[j], acts[j % total])
    barrier()
    def wrapped_step(i):
      j = i % nsets
      sts[j] = wenv.step(sts[j], acts[args.warmup + i])
    ms_wenv = timed_chunks(wrapped_step)
    qps = [st.qp for st in sts]
    # "next" row N2 with the policy inside the launch (bx_policy): the acting loop of training/acting.py:53-74 with
    # PPO's default policy network (ppo/networks.py:67-69: 4 x 32 swish, 2A outputs), sampled actions
    # (caller-supplied N(0,1) noise), observation normalisation on; T = 10 steps per launch, env outputs every step
    from brax_b200.system import Policy
    sizes = (env_obs, 32, 32, 32, 32, 2 * desc.act_size)
    pol = Policy([((torch.rand(sizes[i], sizes[i + 1], device=dev) * 2 - 1) * (3.0 / sizes[i]) ** 0.5,
                   torch.zeros(sizes[i + 1], device=dev)) for i in range(len(sizes) - 1)], activation='swish',
                 obs_mean=torch.zeros(env_obs, device=dev), obs_std=torch.ones(env_obs, device=dev))
    pol_noise = torch.randn(10, b, desc.act_size, device=dev)
    pol_sets = min(nsets, max(2, -(-2 * 126_000_000 // (b * 10 * 4 * (env_obs + 2 * desc.act_size + 8)))))
    pol_obs = [env.sys.observe(q) for q in qps[:pol_sets]]
    n_pol = max(5, args.steps // 10)
    def policy_rollout(i):
      j = i % pol_sets
      qps[j], _, eo = env.sys.rollout(qps[j], None, env=True, policy=pol, obs0=pol_obs[j], noise=pol_noise, steps=10,
                                      info=False)
      pol_obs[j].copy_(eo['obs'][-1])    # (a view would pin the whole [T,B,obs] block: a fresh cudaMalloc per launch)
    for i in range(2 * pol_sets):        # also warms torch's caching allocator: no cudaMalloc in the timed launches
      policy_rollout(i)
    barrier()
    pe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_pol)]
    for i, (x, y) in enumerate(pe):
      x.record()
      policy_rollout(i)
      y.record()
    torch.cuda.synchronize()
    ms_pol = sorted(x.elapsed_time(y) for x, y in pe)[n_pol // 2]
    # "next" row N3: Env.reset on the GPU (noisy joint state -> bx_default_qp -> observe launch)
    for i in range(3):
      env.reset(b, device=dev, seed=i)
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    r0.record()
    for i in range(10):
      env.reset(b, device=dev, seed=i)
    r1.record()
    torch.cuda.synchronize()
    ms_reset = r0.elapsed_time(r1) / 10
  ms_steps = [x.elapsed_time(y) for x, y in ev]
  ms_total = float(sum(ms_steps))
  ms_sorted = sorted(ms_steps)
  ms_fl_steps = sorted(x.elapsed_time(y) for x, y in ev_fl)
  ms_flushed = float(sum(ms_fl_steps))
  ms_b2b = e0.elapsed_time(e1)
  finite = bool(torch.isfinite(qp.pos).all().item())

  # ---- e2e: host buffers through System.step (bx_step_host / bx_step_host_async) ---------------
  # Inputs start in page-locked host memory; every step uploads QP + act, steps, and downloads QP + Info
  # into page-locked host memory (HostBuffers ping-pong so a step's outputs are the next step's inputs,
  # like a host-side rollout loop of the reference's numpy path).
  #   single chain: the whole batch as ONE dependent chain of synchronous calls (round 1's figure): each
  #     step waits for the previous step's last byte, so the link idles while tiles start and drain;
  #   two groups (the e2e figure): the batch as two independent halves, each its own chain, stepped
  #     through System.step_async on two slots -- one half's results travel down while the other half's
  #     inputs travel up and compute.  Same envs, same bytes per step, same public API.
  e2e_steps = 1 if args.no_e2e else max(3, min(args.steps, 50))
  hact = acts[args.warmup:].cpu().pin_memory().numpy()
  h2d, d2h = s.host_transfer_bytes(b)

  def host_chain(lo, hi):
    n_e = hi - lo
    hb = [s.host_buffers(n_e), s.host_buffers(n_e)]
    for dst, src in zip((hb[0].out['pos_o'], hb[0].out['rot_o'], hb[0].out['vel_o'], hb[0].out['ang_o']),
                        (pos, rot, vel, ang)):
      dst[...] = src[lo:hi]
    ha = [np.ascontiguousarray(hact[i % hact.shape[0], lo:hi]) for i in range(min(e2e_steps + 3, hact.shape[0]))]
    return {'hb': hb, 'qp': QP(hb[0].out['pos_o'], hb[0].out['rot_o'], hb[0].out['vel_o'], hb[0].out['ang_o']),
            'acts': ha, 'act_buf': [hb[0].act, hb[1].act]}   # page-locked action buffers (one per in-flight step)

  one = host_chain(0, b)
  for i in range(3):
    one['act_buf'][i % 2][...] = one['acts'][i % len(one['acts'])]
    one['qp'], _ = s.step(one['qp'], one['act_buf'][i % 2], out=one['hb'][(i + 1) % 2])
  barrier()
  t0 = time.perf_counter()
  for i in range(e2e_steps):
    one['act_buf'][i % 2][...] = one['acts'][i % len(one['acts'])]
    one['qp'], _ = s.step(one['qp'], one['act_buf'][i % 2], out=one['hb'][i % 2])
  barrier()
  t_e2e_single = time.perf_counter() - t0
  e2e_finite = bool(np.isfinite(one['qp'].pos).all())
  del one
  halves = [host_chain(0, b // 2), host_chain(b // 2, b)]

  def issue(g, i):
    c = halves[g]
    c['act_buf'][i % 2][...] = c['acts'][i % len(c['acts'])]
    c['pending'] = s.step_async(c['qp'], c['act_buf'][i % 2], out=c['hb'][(i + 1) % 2], slot=g)

  def finish(g):
    s.wait(g)
    halves[g]['qp'] = halves[g]['pending'][0]

  for i in range(3):
    for g in range(2):
      issue(g, i)
    for g in range(2):
      finish(g)
  barrier()
  t0 = time.perf_counter()
  issue(0, 0)
  issue(1, 0)
  for i in range(1, e2e_steps):     # while one half is in flight the host turns the other around
    finish(0)
    issue(0, i)
    finish(1)
    issue(1, i)
  finish(0)
  finish(1)
  barrier()
  t_e2e = time.perf_counter() - t0
  e2e_finite = e2e_finite and all(bool(np.isfinite(c['qp'].pos).all()) for c in halves)
  del halves

  # ---- e2e of the gym-style loop (what envs/to_torch.py moves per step): the state stays on the device,
  # page-locked host ACTIONS in, obs / reward / done / metrics out, one launch working on the mapped host
  # memory; the host waits for every step's results before it issues the next (a policy would need them).
  t_gym, gym_bytes = 0.0, (0, 0)
  if ms_env > 0 and not args.no_e2e:
    hio = env.host_io(b)
    st_g = State(qps[0], None, None, None)
    hact_t = [torch.from_numpy(hact[i % hact.shape[0]]).pin_memory() for i in range(4)]
    for i in range(3):
      st_g = env.step(st_g, hact_t[i % 4], host_out=hio)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
      st_g = env.step(st_g, hact_t[i % 4], host_out=hio)
      torch.cuda.current_stream().synchronize()
    barrier()
    t_gym = time.perf_counter() - t0
    e2e_finite = e2e_finite and bool(torch.isfinite(hio['obs']).all())
    gym_bytes = (4 * b * a, 4 * b * (env_obs + 2 + env.sys.num_metrics))

  # ---- the other BASELINE configs (humanoid 8192, grasp 4096, ant 128 x 1000), each rank its shard ----
  sec = []
  if not args.no_configs and args.env == 'ant' and args.envs == 8192:
    for name, sb, ssteps in SECONDARY:
      sec.append(measure_config(torch, name, sb, ssteps or args.steps, args.warmup, local, rank, barrier))
  # ---- per-rank statistics of the final state: summed over ranks below ----
  mass = torch.tensor([x['mass'] for x in desc.bodies], dtype=torch.float32, device=dev)
  q_fin = qps[0]
  energy = float((0.5 * mass[None, :, None] * q_fin.vel.double() ** 2).sum().item())
  nonfinite = int(sum((~torch.isfinite(getattr(q_fin, k))).sum().item() for k in ('pos', 'rot', 'vel', 'ang')))
  reward_sum = float(last_reward.double().sum().item()) if last_reward is not None else 0.0

  # ---- max over ranks (timings), sum over ranks (statistics) --------------------------
  from brax_b200 import shard
  ms_total, ms_b2b, t_e2e, t_wall, ms_flushed, ms_roll, ms_av, ms_env, ms_chunk = shard.reduce_stats(
      [ms_total, ms_b2b, t_e2e, t_wall, ms_flushed, ms_roll, ms_av, ms_env, ms_chunk], ['max'] * 9, device=dev)
  t_e2e_single, t_gym = shard.reduce_stats([t_e2e_single, t_gym], ['max'] * 2, device=dev)
  sec_ms = shard.reduce_stats([x for m, _, _ in sec for x in m], ['max'] * (3 * len(sec)), device=dev)
  reward_sum, energy, nonfinite, sec_bad = shard.reduce_stats(
      [reward_sum, energy, nonfinite, sum(0 if i['finite'] else 1 for _, i, _ in sec)], ['sum'] * 4, device=dev)
  value = world * b * args.steps / (ms_chunk * 1e-3)
  value_per_step_events = world * b * args.steps / (ms_total * 1e-3)
  value_b2b = world * b * args.steps / (ms_b2b * 1e-3)
  value_flushed = world * b * args.steps / (ms_flushed * 1e-3)
  value_flushed_median = world * b / (ms_fl_steps[len(ms_fl_steps) // 2] * 1e-3)
  e2e_value = world * b * e2e_steps / t_e2e

  if rank == 0:
    hbm_gbs, sm_max_mhz, src = load_peaks()
    clocks = clk.summary()
    flops = load_flops(args.env)
    per_gpu = value / world
    bytes_env = desc.bytes_per_env_step()
    out = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world,
        'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_chunk / args.steps, 'higher_is_better': True,
        'per_step_event_pairs': {'value': value_per_step_events, 'ms_per_step': ms_total / args.steps},
        'ms_per_step_min_median_max': [ms_sorted[0], ms_sorted[len(ms_sorted) // 2], ms_sorted[-1]],
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {
            'workload': workload_name(args, desc),
            'l2': 'inputs larger than L2: the timed steps rotate over %d independent env batches '
                  '(%.0f MB working set vs 126 MB L2), so each step reads state last touched %d '
                  'steps earlier; the timed steps are enqueued in chunks of <= 128 launches behind a spin kernel '
                  'and each chunk is bracketed by ONE event pair (device time only, no host gaps, no '
                  'event records between the launches)' % (nsets, nsets * bytes_set / 1e6, nsets),
            'env_batches': nsets,
            'threads_per_tile': s_threads(s), 'bodies': n, 'contacts': k, 'act': a},
        # one env batch, whole L2 (kernel code and tables included) evicted before every step: the MEDIAN
        # step is the figure; the mean beside it carries the run's rare multi-100-us hiccups of single
        # launches right after a 256 MiB fill (max in ms_l2_flushed_min_median_max)
        'value_l2_flushed': value_flushed_median,
        'value_l2_flushed_mean_step': value_flushed,
        'ms_per_step_l2_flushed': ms_flushed / args.steps,
        'ms_l2_flushed_min_median_max': [ms_fl_steps[0], ms_fl_steps[len(ms_fl_steps) // 2], ms_fl_steps[-1]],
        # a plain Python loop `qp, info = sys.step(qp, act)` with NO spin kernel in front: host enqueue
        # rate included (event pair around the whole loop) ...
        'value_back_to_back': value_b2b,
        # ... and System.make_stepper: 8 in-place steps per CUDA-graph replay, same event method, ONE env batch
        # stepped in place (its state stays in L2, unlike `value`'s rotation over cold batches)
        'value_stepper_graph': world * b / (ms_graph * 1e-3),
        'with_angle_vel_epilogue': {'value': world * b * args.steps / (ms_av * 1e-3), 'unit': UNIT,
                                    'ms_per_step': ms_av / args.steps,
                                    'extra_bytes_per_env_step': 24 * len(desc.joints)},
        'env_step': ({'value': world * b * args.steps / (ms_env * 1e-3), 'unit': UNIT,
                      'ms_per_step': ms_env / args.steps, 'obs_size': env_obs,
                      'what': 'full Env.step (sys.step + obs/reward/done/metrics) in one launch',
                      'extra_bytes_per_env_step': 4 * (env_obs + 12),
                      'with_episode_and_auto_reset_wrappers': {
                          'value': world * b * args.steps / (ms_wenv * 1e-3), 'unit': UNIT,
                          'ms_per_step': ms_wenv / args.steps, 'episode_length': 1000},
                      'policy_rollout': {
                          'value': world * b * 10 / (ms_pol * 1e-3), 'unit': UNIT, 'ms_per_step': ms_pol / 10,
                          'steps_per_launch': 10, 'timing': 'median launch',
                          'what': 'actions = policy(obs); state = env.step(state, actions) x 10 in one launch (bx_policy: '
                                  'normalize -> MLP 4 x 32 swish -> tanh-normal sample), no actions read from memory'},
                      'reset_ms': ms_reset,
                      'reset_what': 'Env.reset of the whole batch through the Python mirror (host-bound at this size): '
                                    'joint noise, bx_default_qp forward kinematics, observe launch'} if ms_env > 0 else None),
        'rollout': {'steps_per_launch': roll_t, 'launches': n_roll,
                    'value': world * b * roll_t * n_roll / (ms_roll * 1e-3), 'unit': UNIT,
                    'ms_per_step': ms_roll / (roll_t * n_roll), 'timing': 'median launch'},
        'wall_s_timed_loop_incl_spin': t_wall,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d,
                'd2h_bytes_per_step': d2h, 'steps': e2e_steps, 'finite': e2e_finite,
                'gbs_per_rank': {'h2d': h2d * e2e_steps / t_e2e / 1e9, 'd2h': d2h * e2e_steps / t_e2e / 1e9},
                'path': 'System.step_async / wait (numpy QP + act on page-locked HostBuffers, QP + full Info out) -> '
                        'bx_step_host_async: the batch as two independent halves on two slots, each half a dependent '
                        'chain step -> step; every launch works on the mapped host memory (tiles pull inputs / push '
                        'outputs over PCIe)',
                'single_chain': {'value': world * b * e2e_steps / t_e2e_single, 'unit': UNIT,
                                 'what': 'the whole batch as ONE dependent chain of synchronous System.step calls '
                                         '(bx_step_host; round 1\'s e2e figure)'}},
        'e2e_env_step': ({'value': world * b * e2e_steps / t_gym, 'unit': UNIT, 'h2d_bytes_per_step': gym_bytes[0],
                          'd2h_bytes_per_step': gym_bytes[1], 'steps': e2e_steps,
                          'gbs_per_rank': {'h2d': gym_bytes[0] * e2e_steps / t_gym / 1e9, 'd2h': gym_bytes[1] * e2e_steps / t_gym / 1e9},
                          'path': 'Env.step(state on the device, page-locked host actions, host_out=page-locked obs / reward / '
                                  'done / metrics): the gym-style loop of brax/envs/wrappers.py:228-260 + to_torch.py:56-64; the '
                                  'host synchronises after every step'} if t_gym > 0 else None),
        'gpu_launches': int(launches),
        'clocks': clocks, 'finite': finite and nonfinite == 0 and sec_bad == 0,
        'stats_all_ranks': {'sum_reward_last_env_step': reward_sum, 'sum_kinetic_energy_final_state': energy,
                            'n_nonfinite': int(nonfinite),
                            'what': 'summed over the %d ranks in the final reduce (nccl); finite covers every rank and config' % world},
    }
    # dominant (only) kernel: bx_step_kernel, one launch per step; its average
    # duration IS ms_per_step (CUDA events bracket each launch on its stream).
    hbm_ach = per_gpu * bytes_env / 1e9
    out['roofline'] = {
        'bound': 'hbm', 'achieved': hbm_ach, 'peak': hbm_gbs, 'unit': 'GB/s',
        'frac': hbm_ach / hbm_gbs, 'traffic': NCU_TRAFFIC.get(args.env, (None, None))[0],
        'traffic_source': NCU_TRAFFIC.get(args.env, (None, 'no ncu --set full capture of this env committed'))[1],
        'peak_source': src, 'bytes_per_env_step': bytes_env,
        'bytes_per_launch': bytes_env * b,
        'note': 'schema roofline (HBM). The kernel is NOT HBM-bound: it is FP32-pipe / '
                'instruction-issue bound (~60 flop/B), see roofline_fp32'}
    if flops:
      mhz = clocks['sm_mhz'] or sm_max_mhz
      peak_max = SM_COUNT * FP32_LANES * 2 * sm_max_mhz * 1e6 / 1e12
      ach = per_gpu * flops['flops'] / 1e12
      out['roofline_fp32'] = {
          'bound': 'fp32', 'achieved': ach, 'peak': peak_max, 'unit': 'TFLOP/s',
          'frac': ach / peak_max, 'flops_per_env_step': flops['flops'],
          'peak_at_observed_clock': SM_COUNT * FP32_LANES * 2 * mhz * 1e6 / 1e12,
          'note': 'flops = exact add+mul+div+sqrt+trig count of the oracle (oracle/flops.json); '
                  'peak = 148 SM x 128 lanes x 2 (FMA) x sm_max_mhz; tensor cores do not apply'}
    # every BASELINE.json config under one key: the headline first, then the secondary ones
    head = {'env': args.env, 'envs_per_gpu': b, 'steps': args.steps, 'workload': workload_name(args, desc),
            'value': value, 'unit': UNIT, 'ms_per_step': ms_chunk / args.steps,
            'roofline': {k: out['roofline'][k] for k in ('bound', 'achieved', 'peak', 'unit', 'frac', 'traffic', 'traffic_source')},
            'roofline_fp32': ({k: out['roofline_fp32'][k] for k in ('bound', 'achieved', 'peak', 'unit', 'frac')} if flops else None),
            'env_step': ({'value': out['env_step']['value'], 'unit': UNIT, 'ms_per_step': out['env_step']['ms_per_step']}
                         if out['env_step'] else None),
            'rollout': {k: out['rollout'][k] for k in ('steps_per_launch', 'value', 'unit', 'ms_per_step', 'timing')}}
    out['configs'] = [head] + [config_entry(info, d2, world, sec_ms[3 * i], sec_ms[3 * i + 1], sec_ms[3 * i + 2],
                                            hbm_gbs, sm_max_mhz) for i, (_, info, d2) in enumerate(sec)]
    if world == 1 and not args.no_cpu:
      out['cpu_baseline'] = cpu_baseline(desc, args.env, min(b, 8192), budget_s=args.cpu_seconds)
    print(json.dumps(out))
  if world > 1:
    dist.destroy_process_group()


def s_threads(s):
  return int(s._lib.bx_system_threads(s._h))


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--gpus', type=int, default=1)
  ap.add_argument('--steps', type=int, default=100)
  ap.add_argument('--warmup', type=int, default=10)
  ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
  ap.add_argument('--env', default='ant')
  ap.add_argument('--envs', type=int, default=8192, help='environments per GPU')
  ap.add_argument('--threads', type=int, default=0, help='threads per 32-env tile (0 = default)')
  ap.add_argument('--contact-section', type=int, default=0,
                  help='0 = default, 1 = four contact phases, 2 = fused per-body section')
  ap.add_argument('--autotune', action='store_true', help='time CTA sizes on the first step')
  ap.add_argument('--cpu-seconds', type=float, default=12.0)
  ap.add_argument('--no-cpu', action='store_true')
  ap.add_argument('--no-e2e', action='store_true')
  ap.add_argument('--sweep', default='', help='comma-separated TOTAL batch sizes (BASELINE config 5), split over the ranks')
  ap.add_argument('--no-configs', action='store_true', help='skip the secondary BASELINE configs')
  args = ap.parse_args()
  from brax_b200.desc import SystemDesc
  desc = SystemDesc.load(os.path.join(ROOT, 'brax_b200', 'systems', args.env + '.json'))
  if args.sweep and args.impl != 'reference':
    run_sweep(args, desc)
  elif args.impl == 'reference':
    run_reference(args, desc)
  else:
    run_gpu(args, desc)


if __name__ == '__main__':
  main()
