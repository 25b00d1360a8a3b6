"""Host-side mirror of the reference's `brax.System` for the step path.

    sys = System(desc)                      # or System.from_brax(brax_system)
    qp, info = sys.step(qp, act)            # brax/physics/system.py:244

`QP`, `P`, `Info` carry the same field names and shapes as the reference's
flax dataclasses (brax/physics/base.py:28-153).  `step` accepts either

  * torch CUDA tensors  -> `bx_step` on torch's current stream (no sync, no
    allocation inside the library; outputs are fresh tensors), or
  * numpy arrays (host) -> `bx_step_host` (H2D, one launch, D2H, sync),

batched `[B, N, ...]` or unbatched `[N, ...]` exactly like `vmap(System.step)`
vs `System.step`.  Only `step` runs on the GPU; `default_qp` returns the state
the reference's `default_qp()` produced when the description was extracted.
Everything is float32.  No CPU fallback: without the CUDA library this raises.
"""
import ctypes as C
import dataclasses
from typing import Any, Tuple

import numpy as np

from brax_b200 import lib as _lib
from brax_b200.desc import POLICY_ACT, POLICY_MAX_LAYERS, CPolicy, CStepIO, SystemDesc


@dataclasses.dataclass
class P:
  """Time derivatives (base.py:50-72)."""
  vel: Any
  ang: Any


@dataclasses.dataclass
class QP:
  """Coordinate and time-derivative frame of every body (base.py:75-133)."""
  pos: Any
  rot: Any
  vel: Any
  ang: Any


@dataclasses.dataclass
class Info:
  """Side outputs of a physics step (base.py:136-153)."""
  contact: P
  joint: P
  actuator: P
  contact_pos: Any
  contact_normal: Any
  contact_penetration: Any


class _LazyInfo(Info):
  """`Info` of a torch-path step whose tensors (views of the step's one output block) are built
  on first access: a stepping loop that never looks at the Info pays nothing for it."""

  def __init__(self, views):
    object.__setattr__(self, '_views', views)   # name -> zero-argument callable returning the view

  def __getattr__(self, name):                  # only reached for fields not built yet
    v = object.__getattribute__(self, '_views')
    if name in ('contact', 'joint', 'actuator'):
      val = P(v[name + '_vel'](), v[name + '_ang']())
    elif name in ('contact_pos', 'contact_normal', 'contact_penetration'):
      val = v[name]()
    else:
      raise AttributeError(name)
    object.__setattr__(self, name, val)
    return val

  def __repr__(self):
    return 'Info(<views of the step output block>)'


_PTR_FIELDS = [n for n, t in CStepIO._fields_ if t is C.c_void_p]
_PTR_INDEX = {n: i for i, n in enumerate(_PTR_FIELDS)}


class _Plan:
  """Per (batch, outputs) call plan of the torch path: its own CStepIO, viewed as an array of
  pointer words so that a call writes all output pointers with one vectorised store."""

  def __init__(self, specs):
    self.names, self.shapes, self.offs, self.counts = [], [], [], []
    off = 0
    for name, shape in specs:
      cnt = 1
      for d in shape:
        cnt *= int(d)
      self.names.append(name)
      self.shapes.append(tuple(shape))
      self.offs.append(off)
      self.counts.append(cnt)
      off += -(-cnt // 4) * 4          # every slice starts on a 16-byte boundary (vector stores)
    self.total = max(off, 4)
    self.io = CStepIO()                 # zero-initialised: every optional pointer NULL
    self.words = np.ctypeslib.as_array((C.c_uint64 * len(_PTR_FIELDS)).from_address(C.addressof(self.io)))
    live = [i for i, c in enumerate(self.counts) if c]
    self.out_idx = np.asarray([_PTR_INDEX[self.names[i]] for i in live], dtype=np.int64)
    self.out_off = np.asarray([4 * self.offs[i] for i in live], dtype=np.uint64)
    self.index = {n: i for i, n in enumerate(self.names)}

  def view(self, flat, name):
    i = self.index[name]
    return flat[self.offs[i]:self.offs[i] + self.counts[i]].view(self.shapes[i])


_OUT = (('pos_o', 3, 'n'), ('rot_o', 4, 'n'), ('vel_o', 3, 'n'),
        ('ang_o', 3, 'n'), ('contact_vel', 3, 'n'), ('contact_ang', 3, 'n'),
        ('joint_vel', 3, 'n'), ('joint_ang', 3, 'n'), ('actuator_vel', 3, 'n'),
        ('actuator_ang', 3, 'n'), ('contact_pos', 3, 'k'),
        ('contact_normal', 3, 'k'), ('contact_penetration', 0, 'k'))
# optional epilogue: Joint.angle_vel of the new state, [B, J, 3] each
_OUT_AV = (('joint_angle', 3, 'j'), ('joint_velocity', 3, 'j'))
# optional env epilogue (systems created with an EnvDesc): Env.step's outputs
_OUT_ENV = (('obs', 1, 'o'), ('reward', 0, 'one'), ('done', 0, 'one'), ('metrics', 1, 'm'))


@dataclasses.dataclass
class JointState:
  """`Joint.angle_vel` of every joint (joints.py:197-226) before the free_dofs
  selection: angle / vel are [B, J, 3] (or [J, 3]), slots beyond a joint's dof are 0.
  `System.angle_vel_groups` turns it into what the reference's envs consume."""
  angle: Any
  vel: Any


@dataclasses.dataclass
class AutoReset:
  """What EpisodeWrapper / AutoResetWrapper keep in `state` (envs/wrappers.py:95-148), handed to
  `System.step(..., env=True, auto_reset=...)` so that both wrappers run inside the launch."""
  first_qp: 'QP'            # state.info['first_qp'] [B,N,*]
  first_obs: Any            # state.info['first_obs'] [B,O]
  done: Any                 # state.done of the state passed in [B]
  steps: Any = None         # state.info['steps'] [B] (None: zeros)
  episode_length: int = 0   # EpisodeWrapper.episode_length; 0 = no EpisodeWrapper
  action_repeat: int = 1    # EpisodeWrapper.action_repeat (wrappers.py:106-111)


@dataclasses.dataclass
class Policy:
  """An MLP policy evaluated INSIDE a rollout launch (C ABI `bx_policy`): the reference's policy network
  (training/networks.py:83-102) on the normalised obs, NormalTanhDistribution's mode or sample
  (training/distribution.py:131-158) as the action.

  params: list of (W [in, out], b [out]) torch CUDA tensors per dense layer (flax `kernel`, `bias`), or one
  flat float32 CUDA tensor already in the ABI's order; the last layer is 2A wide (loc, scale) or A (loc).
  obs_mean / obs_std: [O] CUDA tensors or None.  activation: 'relu' | 'swish' | 'tanh' (hidden layers)."""
  params: Any
  sizes: Any = None          # [O, h0, ..., out]; inferred from a list of (W, b)
  activation: str = 'swish'
  obs_mean: Any = None
  obs_std: Any = None
  min_std: float = 0.001

  def pack(self):
    """-> (CPolicy without noise, tensors to keep alive)."""
    import torch
    if isinstance(self.params, (list, tuple)):
      sizes = [int(self.params[0][0].shape[0])] + [int(w.shape[1]) for w, _ in self.params]
      flat = torch.cat([torch.cat([w.to(torch.float32).contiguous().reshape(-1), b.to(torch.float32).reshape(-1)])
                        for w, b in self.params])
    else:
      sizes, flat = [int(x) for x in self.sizes], self.params.to(torch.float32).contiguous()
    if not 2 <= len(sizes) <= POLICY_MAX_LAYERS + 1:
      raise ValueError('a policy has 1..%d dense layers' % POLICY_MAX_LAYERS)
    if flat.numel() != sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1)):
      raise ValueError('params do not match sizes %s' % sizes)
    if not flat.is_cuda:
      raise RuntimeError('policy parameters must live on the GPU (no CPU fallback)')
    cp = CPolicy()
    cp.struct_size = C.sizeof(CPolicy)
    cp.num_layers = len(sizes) - 1
    for i, v in enumerate(sizes):
      cp.sizes[i] = v
    cp.activation = POLICY_ACT[self.activation]
    cp.min_std = float(self.min_std)
    keep = [flat]
    cp.params = flat.data_ptr()
    for name in ('obs_mean', 'obs_std'):
      t = getattr(self, name)
      if t is not None:
        t = t.to(torch.float32).contiguous()
        if t.numel() != sizes[0] or not t.is_cuda:
          raise ValueError('%s must be a CUDA tensor of %d elements' % (name, sizes[0]))
        keep.append(t)
        setattr(cp, name, t.data_ptr())
    return cp, keep


class HostBuffers:
  """Page-locked host arrays for the host path of `System.step` (bx_host_alloc).

  `out` holds one array per output of the step; `qp()` / `act` are optional
  pinned input arrays a caller can fill in place.  Passing `out=bufs` to
  `System.step` writes the results into these arrays (they are overwritten by
  the next call that uses the same HostBuffers)."""

  def __init__(self, system: 'System', batch: int):
    self._lib = system._lib
    self._ptrs = []
    n, k, a = system.num_bodies, system.num_contacts, system.act_size

    def block(shapes):
      """One page-locked allocation holding arrays of the given shapes back to back (each on a 16-byte
      boundary): the outputs of a step are then one contiguous run of host memory, which
      bx_step_host_async moves with a single DMA."""
      counts = [int(np.prod(sh)) for sh in shapes]
      offs, off = [], 0
      for c in counts:
        offs.append(off)
        off += -(-max(c, 1) // 4) * 4
      p = self._lib.bx_host_alloc(off * 4)
      if not p:
        raise MemoryError('bx_host_alloc: ' + _lib.last_error())
      self._ptrs.append(p)
      whole = np.ctypeslib.as_array((C.c_float * off).from_address(p))
      return [(whole[o:o + c].reshape(sh), p + 4 * o) for o, c, sh in zip(offs, counts, shapes)]

    def arr(shape):
      return block([shape])[0][0]

    self.out = {}
    self.out_ptr = {}   # name -> address (None for empty outputs), read by System.step
    specs = [(name, ((batch, n if dim == 'n' else k, w) if w else (batch, n if dim == 'n' else k))) for name, w, dim in _OUT]
    nj = len(system.desc.joints)
    specs += [(name, (batch, nj, w)) for name, w, _ in _OUT_AV]
    if system.env is not None:
      specs += [('obs', (batch, system.obs_size)), ('reward', (batch,)), ('done', (batch,)),
                ('metrics', (batch, system.num_metrics))]
    for (name, _), (a_, ptr) in zip(specs, block([sh for _, sh in specs])):
      self.out[name] = a_
      self.out_ptr[name] = ptr if a_.size else None
    self.pos, self.rot = arr((batch, n, 3)), arr((batch, n, 4))
    self.vel, self.ang = arr((batch, n, 3)), arr((batch, n, 3))
    self.act = arr((batch, a))
    self.batch = batch

  def qp(self) -> 'QP':
    return QP(self.pos, self.rot, self.vel, self.ang)

  def __del__(self):
    for p in getattr(self, '_ptrs', []):
      try:
        self._lib.bx_host_free(p)
      except Exception:
        pass
    self._ptrs = []


def _is_torch(x) -> bool:
  return type(x).__module__.startswith('torch')


def _ingest(x):
  """Arrays that are neither torch tensors nor numpy: anything that speaks DLPack (`__dlpack__`: JAX
  arrays, CuPy, ...) or the CUDA array interface (`__cuda_array_interface__`: Numba, CuPy) is viewed --
  zero copy -- as a torch CUDA tensor (the DLPack hop brax/io/torch.py:48-53,79-88 makes between JAX and
  torch).  torch tensors and numpy arrays pass through."""
  if _is_torch(x) or isinstance(x, np.ndarray) or isinstance(x, (list, tuple, float, int)):
    return x
  if hasattr(x, '__dlpack__'):
    import torch
    return torch.from_dlpack(x)
  if hasattr(x, '__cuda_array_interface__'):
    import torch
    return torch.as_tensor(x, device=torch.device('cuda', int(getattr(x, 'device_index', 0) or 0)))
  return x


def tile_bytes(desc: SystemDesc, env=None):
  """(shared-memory bytes of one 32-env tile, quad rows per region) of a system
  (with the env epilogue of `env`, an EnvDesc, if given); needs no device
  (bx_desc_tile_bytes)."""
  lib = _lib.load()
  cdesc, keep = desc.to_c(env)
  nbytes = C.c_int64()
  rows = (C.c_int32 * 8)()
  rc = lib.bx_desc_tile_bytes(C.byref(cdesc), C.byref(nbytes), rows)
  del keep
  if rc != 0:
    raise ValueError('bx_desc_tile_bytes: ' + _lib.last_error())
  names = ('bodies', 'act', 'slots', 'contact_rows', 'pre', 'info_contact', 'info_actuator', 'info_joint')
  return int(nbytes.value), dict(zip(names, list(rows)))


class System:
  """A brax v1 system whose `step` is the fused sm_100a kernel."""

  def __init__(self, desc: SystemDesc, device: int = 0, env=None):
    """env: optional `brax_b200.envs.EnvDesc`; the step can then also return what
    the reference's `Env.step` computes after `sys.step` (`step(..., env=True)`)."""
    self.desc = desc
    self.env = env
    self.device = device
    self.num_bodies = desc.num_bodies
    self.num_contacts = desc.num_contacts
    self.act_size = desc.act_size
    self._lib = _lib.load()
    cdesc, keep = desc.to_c(env)
    handle = C.c_void_p()
    rc = self._lib.bx_system_create(C.byref(cdesc), device, C.byref(handle))
    del keep
    if rc != 0:
      raise RuntimeError('bx_system_create: ' + _lib.last_error())
    self._h = handle
    self.threads_per_tile = 0  # 0 = library default
    self._tuned = set()        # batch sizes bx_system_autotune has seen
    # opt-in: time the candidate CTA sizes on the first big step (bx_system_autotune);
    # the per-system default is calibrated against sweeps (profiles/r01_sweeps.txt) and
    # a one-shot timing on a cold device picked worse configurations in practice
    self.autotune = False
    self._plans = {}           # batch -> output slicing plan
    self._io = CStepIO()
    self.obs_size = self._lib.bx_system_obs_size(self._h) if env is not None else 0
    self.num_metrics = self._lib.bx_system_num_metrics(self._h) if env is not None else 0
    if env is not None and self.obs_size != env.obs_size(desc):
      raise RuntimeError('obs size: library %d, EnvDesc %d' % (self.obs_size, env.obs_size(desc)))

  @classmethod
  def from_brax(cls, brax_system, name: str = 'system', act_size=None,
                device: int = 0) -> 'System':
    """Builds from a constructed reference `brax.System` (the drop-in route)."""
    from brax_b200.extract import describe
    return cls(describe(brax_system, name, act_size=act_size), device)

  def __del__(self):
    h, self._h = getattr(self, '_h', None), None
    if h:
      try:
        self._lib.bx_system_destroy(h)
      except Exception:  # interpreter shutdown
        pass

  @property
  def handle(self) -> int:
    """The `bx_system*` as an integer (FFI attribute)."""
    return int(self._h.value)

  def configure(self, threads_per_block: int = 0, contact_section: int = 0):
    """Tuning knobs (0 = keep): warps per 32-env tile = threads_per_block / 32;
    contact_section 1 = four contact phases, 2 = fused per-body section."""
    if self._lib.bx_system_configure(self._h, threads_per_block, contact_section) != 0:
      raise ValueError(_lib.last_error())
    if threads_per_block:
      self.threads_per_tile = threads_per_block
      self.autotune = False

  # ---- state helpers ---------------------------------------------------------
  def default_angle(self):
    """`System.default_angle()` (system.py:86-110), float32 [D]."""
    return np.asarray(self.desc.reset['default_angle'], dtype=np.float32)

  @property
  def num_joint_dof(self) -> int:
    """D: length of the joint_angle / joint_velocity vectors `default_qp` takes."""
    return int(self.desc.reset['num_dof'])

  def default_qp(self, batch: int = None, device=None, joint_angle=None, joint_velocity=None, joint_offset=None) -> QP:
    """`System.default_qp(joint_angle=..., joint_velocity=...)` (system.py:112-242).

    Without joint arguments: the default state as recorded at extraction time (numpy, or
    torch on `device`).  With joint_angle / joint_velocity ([D] or [B,D] torch CUDA tensors;
    a missing one is `default_angle()` / zeros): forward kinematics and the lowering of free
    trees onto the xy plane in ONE kernel launch (bx_default_qp) -- what every env's reset
    calls with its noisy joint state (envs/ant.py:198-204)."""
    if joint_angle is not None or joint_velocity is not None:
      # joint_offset: per-env joint sockets [B,J,3] (see `step`): the randomised Systems' own default pose
      return self._default_qp_torch(joint_angle, joint_velocity, joint_offset)
    if joint_offset is not None:
      raise ValueError('joint_offset needs joint_angle / joint_velocity (the torch path)')
    d = self.desc.default_qp
    arrs = [np.asarray(d[k], dtype=np.float32) for k in ('pos', 'rot', 'vel', 'ang')]
    if batch is not None:
      arrs = [np.ascontiguousarray(np.broadcast_to(a, (batch,) + a.shape)) for a in arrs]
    if device is not None:
      import torch
      arrs = [torch.from_numpy(a).to(device) for a in arrs]
    return QP(*arrs)

  def joint_offsets(self):
    """`Joint.parent_offset` of every joint as the descriptor holds it, float32 [J,3]: the base to which a joint
    randomiser adds its per-env deltas (`joint_offset=` of `step` / `rollout`)."""
    return np.asarray([j['off_p'] for j in self.desc.joints], np.float32).reshape(-1, 3)

  def observe(self, qp: QP, friction=None):
    """obs [B,O] of a state as `Env.reset` builds it: `_get_obs(qp, info)` with a zero Info and a
    zero action (envs/ant.py:203-204, humanoid.py:230-231, fetch.py:41-42, grasp.py:60-61) -- the
    env epilogue of the step kernel without its substeps, one launch.  torch CUDA tensors;
    systems built with an EnvDesc."""
    import torch
    if self.env is None:
      raise ValueError('this System was built without an EnvDesc')
    if not _is_torch(qp.pos):
      raise RuntimeError('observe takes torch CUDA tensors (no CPU fallback)')
    b = int(qp.pos.shape[0]) if qp.pos.dim() == 3 else 1
    shape = (b, self.act_size) if qp.pos.dim() == 3 else (self.act_size,)
    act = torch.zeros(shape, dtype=torch.float32, device=qp.pos.device)
    return self._step_torch(qp, act, 1, False, True, None, None, False, observe=True, friction=friction)[2]['obs']

  def _default_qp_torch(self, joint_angle, joint_velocity, joint_offset=None) -> QP:
    import torch
    if not self.desc.reset:
      raise ValueError('this descriptor carries no reset tables (SystemDesc.reset)')
    given = joint_angle if joint_angle is not None else joint_velocity
    if not _is_torch(given) or given.device.type != 'cuda':
      raise RuntimeError('joint_angle / joint_velocity must be torch CUDA tensors (no CPU fallback)')
    dev, nd = given.device, self.num_joint_dof
    batched = given.dim() == 2
    b = int(given.shape[0]) if batched else 1
    f32 = torch.float32
    if joint_angle is None:
      joint_angle = torch.from_numpy(self.default_angle()).to(dev).expand(b, nd)
    if joint_velocity is None:
      joint_velocity = torch.zeros(b, nd, dtype=f32, device=dev)
    ja = joint_angle.to(f32).reshape(b, nd).contiguous()
    jv = joint_velocity.to(f32).reshape(b, nd).contiguous()
    n = self.num_bodies
    flat = torch.empty(b * n * 13, dtype=f32, device=dev)
    outs = [flat[:b * n * 3].view(b, n, 3), flat[b * n * 3:b * n * 7].view(b, n, 4),
            flat[b * n * 7:b * n * 10].view(b, n, 3), flat[b * n * 10:].view(b, n, 3)]
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    jo = None
    if joint_offset is not None:
      # per-env joint sockets [B,J,3] (`joint_offset=` of `step`): the pose the randomised System's default_qp builds
      jo = joint_offset.to(f32).reshape(b, len(self.desc.joints), 3).contiguous()
      if jo.device != dev:
        raise ValueError('joint_offset must live on the device of joint_angle')
    rc = self._lib.bx_default_qp_offsets(self._h, b, ja.data_ptr() if nd else None, jv.data_ptr() if nd else None,
                                         None if jo is None else jo.data_ptr(), *[o.data_ptr() for o in outs], stream)
    if rc != 0:
      raise RuntimeError('bx_default_qp: ' + _lib.last_error())
    if not batched:
      outs = [o[0] for o in outs]
    return QP(*outs)

  def make_stepper(self, batch: int, steps_per_replay: int = 8, graph: bool = True, info: bool = False) -> 'Stepper':
    """A stepping loop without per-step host work: `steps_per_replay` chained `System.step`s on
    static device buffers, captured once into a CUDA graph and replayed (one graph launch per k
    steps).  The reference shape: `jax.jit(step)` is one dispatch per call (envs/wrappers.py:243)."""
    return Stepper(self, batch, steps_per_replay, graph, info)

  def make_state(self, batch: int) -> 'DeviceState':
    """An opaque device-resident state of `batch` environments, stepped in place (C ABI `bx_state`)."""
    return DeviceState(self, batch)

  # ---- the step ---------------------------------------------------------------
  def step(self, qp: QP, act, out: 'HostBuffers' = None, angle_vel: bool = False, env: bool = False,
           env_rand=None, auto_reset: 'AutoReset' = None, info: bool = True, env_out=None, friction=None,
           joint_offset=None):
    """`System.step(qp, act) -> (qp, info)` (system.py:244-247).

    out: optional `HostBuffers` (host path only): results are written into its
    page-locked arrays instead of freshly allocated pageable ones.
    angle_vel: also return `Joint.angle_vel` of the new state as a further value
    (`JointState`), computed in the same launch -- what every env's `_get_obs`
    asks for right after the step (envs/ant.py:259, envs/humanoid.py:285).
    env: (systems built with an EnvDesc) also return, last, a dict with what the
    reference's `Env.step` computes after `sys.step` -- obs [B,O], reward [B],
    done [B], metrics [B,M] (envs/ant.py:222-281, envs/humanoid.py:246-340).
    auto_reset: (with env=True, torch path) an `AutoReset`: EpisodeWrapper.step and
    AutoResetWrapper.step (envs/wrappers.py:105-148) applied inside the launch; the env
    dict gains 'steps' and 'truncation' and qp / obs / done are the wrapped ones.
    info: False (torch path) = do not write the Info outputs (the returned info is None): what
    `Env.step` needs, which drops the Info after building the obs (envs/ant.py:224).
    env_out: (torch path, env=True) dict of preallocated float32 tensors 'obs' [B,O], 'reward' [B],
    'done' [B], 'metrics' [B,M] that receive the env outputs -- CUDA tensors, or PAGE-LOCKED host
    tensors (`pin_memory=True`): the launch then writes them over PCIe itself.  `act` may likewise be a
    page-locked host tensor.  That is the gym-style loop with the state resident on the device
    (brax/envs/wrappers.py:228-260 + to_torch.py:56-64): actions in, obs / reward / done out, one
    launch, no separate copies; synchronise the stream before reading the host tensors.
    friction: (torch path) float32 CUDA tensor [B]: per-env collider material friction -- the domain
    randomisation of experimental/tracing/randomizers.py:26-51 (every collider of env e has
    material.friction = friction[e]; pair friction = friction[e]^2).
    joint_offset: (torch path) float32 CUDA tensor [B,J,3]: per-env `Joint.parent_offset` of every joint -- the
    joint randomiser of experimental/tracing/randomizers.py:54-97 (`joint_offsets()` + your per-env deltas)."""
    if env and self.env is None:
      raise ValueError('this System was built without an EnvDesc')
    if auto_reset is not None and not env:
      raise ValueError('auto_reset needs env=True')
    if not (_is_torch(qp.pos) or isinstance(qp.pos, np.ndarray)):
      # DLPack / CUDA-array-interface producers (JAX, CuPy, Numba): zero-copy torch views
      qp = QP(_ingest(qp.pos), _ingest(qp.rot), _ingest(qp.vel), _ingest(qp.ang))
      act = _ingest(act)
    elif _is_torch(qp.pos) and not _is_torch(act):
      act = _ingest(act)
      if isinstance(act, np.ndarray):
        raise TypeError('qp is on the GPU, act is a numpy array: pass a CUDA (or page-locked) tensor')
    if _is_torch(qp.pos):
      return self._step_torch(qp, act, 1, angle_vel, env, env_rand, auto_reset, info, env_out=env_out, friction=friction,
                              joint_offset=joint_offset)
    if env_out is not None or friction is not None or joint_offset is not None:
      raise ValueError('env_out / friction / joint_offset belong to the torch path')
    if auto_reset is not None:
      raise ValueError('auto_reset is supported with torch CUDA tensors only')
    return self._step_numpy(qp, act, out, angle_vel, env, env_rand)

  def angle_vel_groups(self, js: JointState):
    """[(angle, vel) per joint group] exactly as `[j.angle_vel(qp) for j in
    sys.joints]`: per group the joints' first free_dofs dofs, flattened
    (joints.py:216-225)."""
    groups = {}
    for i, j in enumerate(self.desc.joints):
      groups.setdefault(j['group'], []).append((i, int(j.get('free_dofs', j['dof']))))
    cat = np.concatenate
    if _is_torch(js.angle):
      import torch
      cat = lambda xs, axis: torch.cat(xs, dim=axis)  # noqa: E731
    res = []
    for g in sorted(groups):
      res.append(tuple(cat([x[..., i, :f] for i, f in groups[g]], -1) for x in (js.angle, js.vel)))
    return res

  def host_buffers(self, batch: int) -> HostBuffers:
    """Page-locked arrays for `step(..., out=...)` at this batch size."""
    return HostBuffers(self, batch)

  def host_transfer_bytes(self, batch: int):
    """(h2d, d2h) bytes one host-path step moves over PCIe."""
    h2d, d2h = C.c_int64(), C.c_int64()
    self._lib.bx_host_transfer_bytes(self._h, batch, C.byref(h2d), C.byref(d2h))
    return int(h2d.value), int(d2h.value)

  def rollout(self, qp: QP, acts, env: bool = False, env_rand=None, auto_reset: 'AutoReset' = None,
              info: bool = True, policy: 'Policy' = None, obs0=None, noise=None, steps: int = None, friction=None,
              joint_offset=None):
    """T steps in ONE launch; acts is [T, B, A] (torch CUDA).  Returns the final
    QP and the Info of the last step (lax.scan of System.step over actions); with
    env=True also the env epilogue of EVERY step: obs [T,B,O], reward / done [T,B],
    metrics [T,B,M]; with auto_reset (see `step`) episodes end and restart inside the
    launch: 'truncation' is [T,B], 'steps' [B] the counters after the last step."""
    if env and self.env is None:
      raise ValueError('this System was built without an EnvDesc')
    if auto_reset is not None and not env:
      raise ValueError('auto_reset needs env=True')
    if policy is not None:
      # acts=None: the actions of every step come from `policy` evaluated inside the launch on the obs of
      # the previous step (obs0 [B,O] = state.obs for the first); the env dict gains 'action' [T,B,A].
      # noise: optional standard normal draws [T,B,A] (stochastic actions); steps: T.
      if not env or obs0 is None or steps is None:
        raise ValueError('a policy rollout needs env=True, obs0 and steps')
      return self._step_torch(qp, None, int(steps), False, True, env_rand, auto_reset, info,
                              policy=(policy, obs0, noise), friction=friction, joint_offset=joint_offset)
    return self._step_torch(qp, acts, int(acts.shape[0]), False, env, env_rand, auto_reset, info, friction=friction,
                            joint_offset=joint_offset)

  def _shapes(self, pos_shape):
    n = self.num_bodies
    if len(pos_shape) == 2:
      batched, b = False, 1
    elif len(pos_shape) == 3:
      batched, b = True, int(pos_shape[0])
    else:
      raise ValueError('qp.pos must be [N,3] or [B,N,3]')
    if int(pos_shape[-2]) != n or int(pos_shape[-1]) != 3:
      raise ValueError('qp.pos has shape %s, system has %d bodies' % (tuple(pos_shape), n))
    return batched, b

  def _specs(self, b, av, env, steps=1, info=True):
    """[(name, shape)] of the outputs of one call."""
    dims = {'n': self.num_bodies, 'k': self.num_contacts, 'j': len(self.desc.joints)}
    base = _OUT if info else _OUT[:4]
    specs = [(name, (b, dims[dim], w) if w else (b, dims[dim])) for name, w, dim in (base + _OUT_AV if av else base)]
    if env:
      lead = (steps, b) if steps > 1 else (b,)
      specs += [('obs', lead + (self.obs_size,)), ('reward', lead), ('done', lead),
                ('metrics', lead + (self.num_metrics,))]
    return specs

  def _wrap(self, out, batched, steps=1):
    env_out = None
    if 'obs' in out:
      env_out = {k: out.pop(k) for k in ('obs', 'reward', 'done', 'metrics', 'steps', 'truncation') if k in out}
      if not batched:
        env_out = {k: (v[:, 0] if steps > 1 and v.ndim > 1 and k != 'steps' else v[0]) for k, v in env_out.items()}
    if not batched:
      out = {k: v[0] for k, v in out.items()}
    qp = QP(out['pos_o'], out['rot_o'], out['vel_o'], out['ang_o'])
    info = None
    if 'contact_vel' in out:
      info = Info(P(out['contact_vel'], out['contact_ang']),
                  P(out['joint_vel'], out['joint_ang']),
                  P(out['actuator_vel'], out['actuator_ang']),
                  out['contact_pos'], out['contact_normal'],
                  out['contact_penetration'])
    res = (qp, info)
    if 'joint_angle' in out:
      res += (JointState(out['joint_angle'], out['joint_velocity']),)
    if env_out is not None:
      res += (env_out,)
    return res

  def _clear_optional(self, io):
    for name, _, _ in _OUT[4:]:
      setattr(io, name, None)
    io.joint_angle = io.joint_velocity = None
    io.obs = io.reward = io.done = io.metrics = None
    io.env_rand = None
    io.first_pos = io.first_rot = io.first_vel = io.first_ang = io.first_obs = None
    io.done_in = io.steps_in = io.steps_out = io.truncation = None
    io.episode_length = 0.0
    io.observe = 0
    io.action_repeat = 0

  def _step_torch(self, qp: QP, act, steps: int, av: bool = False, env: bool = False, env_rand=None,
                  auto_reset=None, info: bool = True, observe: bool = False, env_out=None, policy=None, friction=None,
                  joint_offset=None):
    """env_rand: (fetch / grasp) uniforms of the target teleport, [B,3] or [T,B,3].

    Host cost per call is what bounds a plain Python stepping loop (the kernel takes ~57 us for
    8192 ants), so the call keeps to: five input checks, ONE output allocation, one vectorised store
    of the output pointers into the plan's CStepIO, the C call, four views for the QP; the Info views
    are built lazily (`_LazyInfo`)."""
    import torch
    batched, b = self._shapes(qp.pos.shape)
    n, a = self.num_bodies, self.act_size
    dev = qp.pos.device
    if dev.type != 'cuda':
      raise RuntimeError('torch tensors must live on a CUDA device (no CPU fallback)')
    if dev.index is not None and dev.index != self.device:
      raise RuntimeError('tensors are on %s, system is on cuda:%d' % (dev, self.device))
    f32 = torch.float32

    def prep(t, numel):
      # the common case (float32, contiguous, right size) costs two attribute reads
      if t.dtype is not f32:
        t = t.to(f32)
      if not t.is_contiguous():
        t = t.contiguous()
      if t.numel() != numel:
        raise ValueError('tensor has %d elements, expected %d' % (t.numel(), numel))
      return t

    ins = (prep(qp.pos, b * n * 3), prep(qp.rot, b * n * 4), prep(qp.vel, b * n * 3),
           prep(qp.ang, b * n * 3), prep(act, steps * b * a) if policy is None else qp.pos)
    if ins[4].device.type != 'cuda' and not ins[4].is_pinned():
      raise RuntimeError('host actions must be page-locked (pin_memory): the launch reads them over PCIe')
    ext = env_out is not None
    if ext and not (env and batched and steps == 1):
      raise ValueError('env_out needs env=True on a batched single step')
    key = (b, av, env, steps, info, ext)
    plan = self._plans.get(key)
    if plan is None:
      plan = self._plans[key] = _Plan(self._specs(b, av, env and not ext, steps, info))
    flat = torch.empty(plan.total, dtype=f32, device=dev)
    w = plan.words
    w[:5] = [t.data_ptr() for t in ins]
    w[plan.out_idx] = plan.out_off + np.uint64(flat.data_ptr())
    io = plan.io
    keep = None
    if env_rand is not None:
      rep = int(auto_reset.action_repeat or 1) if auto_reset is not None else 1
      env_rand = prep(env_rand, steps * rep * b * 3)
      io.env_rand = env_rand.data_ptr()
    else:
      io.env_rand = None
    extra = None
    if auto_reset is not None:
      ar = auto_reset
      keep = (prep(ar.first_qp.pos, b * n * 3), prep(ar.first_qp.rot, b * n * 4), prep(ar.first_qp.vel, b * n * 3),
              prep(ar.first_qp.ang, b * n * 3), prep(ar.first_obs, b * self.obs_size), prep(ar.done, b),
              None if ar.steps is None else prep(ar.steps, b))
      io.first_pos, io.first_rot, io.first_vel, io.first_ang, io.first_obs, io.done_in = [t.data_ptr() for t in keep[:6]]
      io.steps_in = None if keep[6] is None else keep[6].data_ptr()
      extra = torch.empty((steps + 1) * b, dtype=f32, device=dev)
      io.steps_out = extra.data_ptr()
      io.truncation = extra.data_ptr() + 4 * b
      io.episode_length = float(ar.episode_length or 0)
      io.action_repeat = int(ar.action_repeat or 1)
    elif io.first_pos:
      io.first_pos = io.first_rot = io.first_vel = io.first_ang = io.first_obs = None
      io.done_in = io.steps_in = io.steps_out = io.truncation = None
      io.episode_length = 0.0
      io.action_repeat = 0
    io.observe = int(observe)
    pol_keep = None
    if policy is not None:
      pol, obs0, noise = policy
      cp, pol_keep = pol.pack()
      obs0 = prep(obs0, b * self.obs_size)
      act_out = torch.empty((steps, b, a) if steps > 1 else (b, a), dtype=f32, device=dev)
      if noise is not None:
        noise = prep(noise, steps * b * a)
        cp.noise = noise.data_ptr()
      pol_keep += [obs0, noise, act_out, cp]
      w[4] = 0                      # act: not read
      io.policy = C.addressof(cp)
      io.obs0 = obs0.data_ptr()
      io.act_out = act_out.data_ptr()
    elif io.policy:
      io.policy = io.obs0 = io.act_out = None
    if friction is not None:
      if not (_is_torch(friction) and friction.is_cuda):
        raise ValueError('friction must be a CUDA tensor [B]')
      friction = prep(friction, b)
      io.friction = friction.data_ptr()
    elif io.friction:
      io.friction = None
    if joint_offset is not None:
      if not (_is_torch(joint_offset) and joint_offset.is_cuda):
        raise ValueError('joint_offset must be a CUDA tensor [B,J,3]')
      joint_offset = prep(joint_offset, b * len(self.desc.joints) * 3)
      io.joint_offset = joint_offset.data_ptr()
    elif io.joint_offset:
      io.joint_offset = None
    if ext:
      for nm, cnt in (('obs', b * self.obs_size), ('reward', b), ('done', b), ('metrics', b * self.num_metrics)):
        t = env_out[nm]
        if t.dtype is not f32 or not t.is_contiguous() or t.numel() != cnt or not (t.is_cuda or t.is_pinned()):
          raise ValueError('env_out[%r] must be a contiguous float32 CUDA or page-locked tensor of %d elements' % (nm, cnt))
        setattr(io, nm, t.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    if self.autotune and not observe and steps == 1 and b >= 1024 and b not in self._tuned:
      # first call at this batch size: time the candidate CTA sizes once
      self._tuned.add(b)
      if self._lib.bx_system_autotune(self._h, b, C.byref(io), stream) != 0:
        raise RuntimeError('bx_system_autotune: ' + _lib.last_error())
    if steps > 1:
      rc = self._lib.bx_rollout(self._h, b, steps, C.byref(io), stream)
    else:
      rc = self._lib.bx_step(self._h, b, C.byref(io), stream)
    if rc != 0:
      raise RuntimeError('bx_step: ' + _lib.last_error())
    if not batched:
      # unbatched call ([N,..] in, [N,..] out): the general wrapper
      out = {name: plan.view(flat, name) for name in plan.names}
      if extra is not None:
        out['steps'] = extra[:b].view((b,))
        out['truncation'] = extra[b:].view((steps, b) if steps > 1 else (b,))
      return self._wrap(out, batched, steps)
    view = plan.view
    res = (QP(view(flat, 'pos_o'), view(flat, 'rot_o'), view(flat, 'vel_o'), view(flat, 'ang_o')),
           _LazyInfo({nm: (lambda nm=nm: view(flat, nm)) for nm in plan.names[4:13]}) if info else None)
    if av:
      res += (JointState(view(flat, 'joint_angle'), view(flat, 'joint_velocity')),)
    if env:
      eo = dict(env_out) if ext else {k: view(flat, k) for k in ('obs', 'reward', 'done', 'metrics')}
      if policy is not None:
        eo['action'] = pol_keep[-2]
      if extra is not None:
        eo['steps'] = extra[:b].view((b,))
        eo['truncation'] = extra[b:].view((steps, b) if steps > 1 else (b,))
      res += (eo,)
    return res

  def step_async(self, qp: QP, act, out: 'HostBuffers', slot: int = 0, angle_vel: bool = False, env: bool = False,
                 env_rand=None):
    """Host path without the wait: page-locked inputs and outputs (`HostBuffers`), the zero-copy launch
    goes to the library's internal stream `slot` (0..2) and the call returns; `wait(slot)` blocks until
    the results are in `out`.  Independent env groups on different slots overlap transfers and compute."""
    return self._step_numpy(qp, act, out, angle_vel, env, env_rand, slot=int(slot))

  def wait(self, slot: int = 0):
    if self._lib.bx_host_wait(self._h, int(slot)) != 0:
      raise RuntimeError('bx_host_wait: ' + _lib.last_error())

  def _step_numpy(self, qp: QP, act, bufs: 'HostBuffers' = None, av: bool = False, env: bool = False,
                  env_rand=None, slot: int = None):
    batched, b = self._shapes(np.shape(qp.pos))
    n, a = self.num_bodies, self.act_size

    def prep(x, numel):
      # float32 C-contiguous arrays of the right size pass through untouched
      if not (isinstance(x, np.ndarray) and x.dtype == np.float32 and x.flags.c_contiguous):
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
      if x.size != numel:
        raise ValueError('array has %d elements, expected %d' % (x.size, numel))
      return x

    ins = (prep(qp.pos, b * n * 3), prep(qp.rot, b * n * 4), prep(qp.vel, b * n * 3),
           prep(qp.ang, b * n * 3), prep(act, b * a))
    io = self._io if slot is None else CStepIO()   # (an in-flight call keeps its own struct)
    io.pos, io.rot, io.vel, io.ang, io.act = [t.__array_interface__['data'][0] for t in ins]
    self._clear_optional(io)
    if bufs is not None:
      if bufs.batch != b:
        raise ValueError('HostBuffers are for batch %d, step has %d' % (bufs.batch, b))
      want = set(name for name, _ in self._specs(b, av, env))
      out = {k: v for k, v in bufs.out.items() if k in want}
      for name in out:
        setattr(io, name, bufs.out_ptr[name])
    else:
      out = {}
      for name, shape in self._specs(b, av, env):
        out[name] = np.empty(shape, dtype=np.float32)
        setattr(io, name, out[name].__array_interface__['data'][0] if out[name].size else None)
    if env_rand is not None:
      env_rand = prep(env_rand, b * 3)
      io.env_rand = env_rand.__array_interface__['data'][0]
    if slot is None:
      rc = self._lib.bx_step_host(self._h, b, C.byref(io))
    else:
      if bufs is None:
        raise ValueError('step_async needs out=HostBuffers (page-locked outputs)')
      rc = self._lib.bx_step_host_async(self._h, b, C.byref(io), slot)
    if rc != 0:
      raise RuntimeError('bx_step_host: ' + _lib.last_error())
    return self._wrap(out, batched)


class DeviceState:
  """Opaque device-resident state of a batch (`System.make_state`, C ABI `bx_state`): the QP lives
  inside the library and is stepped in place; only actions go in and env outputs / Info come out.

      st = sys.make_state(batch); st.set(qp)           # qp: torch CUDA tensors [B,N,...]
      eo = st.rollout(acts, env=True)                   # T steps, ONE launch: obs [T,B,O], reward [T,B] ...
      qp = st.get()                                     # a copy of the current QP

  The carry of `lax.scan(env.step)` (training/acting.py:53-74) that never crosses the boundary."""

  def __init__(self, system: 'System', batch: int):
    self.sys, self.batch = system, int(batch)
    h = C.c_void_p()
    if system._lib.bx_state_create(system._h, self.batch, C.byref(h)) != 0:
      raise RuntimeError('bx_state_create: ' + _lib.last_error())
    self._h = h
    self._plans = {}

  def __del__(self):
    h, self._h = getattr(self, '_h', None), None
    if h:
      try:
        self.sys._lib.bx_state_destroy(h)
      except Exception:
        pass

  def _stream(self):
    import torch
    return C.c_void_p(torch.cuda.current_stream(torch.device('cuda', self.sys.device)).cuda_stream)

  def set(self, qp: QP):
    import torch
    n, b = self.sys.num_bodies, self.batch
    ts = []
    for t, w in ((qp.pos, 3), (qp.rot, 4), (qp.vel, 3), (qp.ang, 3)):
      t = t.to(torch.float32).contiguous()
      if not t.is_cuda or t.numel() != b * n * w:
        raise ValueError('set() takes CUDA tensors [B,N,%d] of this state\'s batch' % w)
      ts.append(t)
    if self.sys._lib.bx_state_set(self._h, *[C.c_void_p(t.data_ptr()) for t in ts], self._stream()) != 0:
      raise RuntimeError('bx_state_set: ' + _lib.last_error())
    return self

  def get(self) -> QP:
    import torch
    n, b = self.sys.num_bodies, self.batch
    dev = torch.device('cuda', self.sys.device)
    out = [torch.empty((b, n, w), dtype=torch.float32, device=dev) for w in (3, 4, 3, 3)]
    if self.sys._lib.bx_state_get(self._h, *[C.c_void_p(t.data_ptr()) for t in out], self._stream()) != 0:
      raise RuntimeError('bx_state_get: ' + _lib.last_error())
    return QP(*out)

  def rollout(self, acts, env: bool = False, info: bool = False, env_rand=None):
    """acts [T,B,A] (or [B,A] for one step), torch CUDA.  Returns (info of the last step or None,
    env outputs of every step or None)."""
    import torch
    sysm, b = self.sys, self.batch
    if env and sysm.env is None:
      raise ValueError('this System was built without an EnvDesc')
    steps = int(acts.shape[0]) if acts.dim() == 3 else 1
    acts = acts.to(torch.float32).contiguous()
    if not acts.is_cuda or acts.numel() != steps * b * sysm.act_size:
      raise ValueError('acts must be a CUDA tensor [T,%d,%d]' % (b, sysm.act_size))
    key = (env, steps, info)
    plan = self._plans.get(key)
    if plan is None:
      plan = self._plans[key] = _Plan(sysm._specs(b, False, env, steps, True)[4:] if info
                                      else sysm._specs(b, False, env, steps, False)[4:])
    flat = torch.empty(plan.total, dtype=torch.float32, device=acts.device)
    plan.words[4] = acts.data_ptr()
    plan.words[plan.out_idx] = plan.out_off + np.uint64(flat.data_ptr())
    io = plan.io
    if env_rand is not None:
      env_rand = env_rand.to(torch.float32).contiguous()
      io.env_rand = env_rand.data_ptr()
    else:
      io.env_rand = None
    if sysm._lib.bx_state_rollout(self._h, steps, C.byref(io), self._stream()) != 0:
      raise RuntimeError('bx_state_rollout: ' + _lib.last_error())
    view = plan.view
    inf = _LazyInfo({nm: (lambda nm=nm: view(flat, nm)) for nm in plan.names[:9]}) if info else None
    eo = {k: view(flat, k) for k in ('obs', 'reward', 'done', 'metrics')} if env else None
    return inf, eo


class Stepper:
  """k chained steps on static buffers, replayed as one CUDA graph (`System.make_stepper`).

      st = sys.make_stepper(batch, steps_per_replay=8)
      st.qp.pos.copy_(pos) ...                 # the state lives in st.qp (static tensors)
      st.acts.copy_(actions)                   # [k, B, A]: step i of a replay reads acts[i]
      st.replay()                              # k steps, in place: st.qp now holds the state k steps on
      st.info                                  # (info=True) the Info of the last step, static tensors

  The launches step IN PLACE (outputs = inputs): a tile reads the whole state of its 32 environments
  into shared memory before it writes anything, and tiles own disjoint environments, so `bx_step`
  allows pos_o == pos; consecutive launches chain through programmatic dependent launch.
  graph=False issues the same k launches directly (for A/B and debugging)."""

  def __init__(self, system: 'System', batch: int, k: int, graph: bool, info: bool):
    import torch
    if k < 1:
      raise ValueError('steps_per_replay must be >= 1')
    self.sys, self.batch, self.k = system, int(batch), int(k)
    dev = torch.device('cuda', system.device)
    n, a, kk = system.num_bodies, system.act_size, system.num_contacts
    f32 = torch.float32
    q0 = system.default_qp(batch=batch)
    self.qp = QP(*[torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in (q0.pos, q0.rot, q0.vel, q0.ang)])
    self.acts = torch.zeros((k, batch, a), dtype=f32, device=dev)
    self.info = None
    self._ios = []
    out = {}
    if info:
      dims = {'n': n, 'k': kk}
      for name, w, dim in _OUT[4:]:
        out[name] = torch.zeros((batch, dims[dim], w) if w else (batch, dims[dim]), dtype=f32, device=dev)
      self.info = Info(P(out['contact_vel'], out['contact_ang']), P(out['joint_vel'], out['joint_ang']),
                       P(out['actuator_vel'], out['actuator_ang']), out['contact_pos'], out['contact_normal'],
                       out['contact_penetration'])
    for i in range(k):
      io = CStepIO()
      io.pos, io.rot, io.vel, io.ang = [t.data_ptr() for t in (self.qp.pos, self.qp.rot, self.qp.vel, self.qp.ang)]
      io.pos_o, io.rot_o, io.vel_o, io.ang_o = io.pos, io.rot, io.vel, io.ang      # in place
      io.act = self.acts[i].data_ptr()
      if info and i == k - 1:
        for name, t in out.items():
          setattr(io, name, t.data_ptr() if t.numel() else None)
      self._ios.append(io)
    self._graph = None
    if graph:
      side = torch.cuda.Stream(device=dev)
      side.wait_stream(torch.cuda.current_stream(dev))
      with torch.cuda.stream(side):          # warm-up on the capture stream (module load, attributes)
        keep = [t.clone() for t in (self.qp.pos, self.qp.rot, self.qp.vel, self.qp.ang)]
        self._launch_all()
        for t, kp in zip((self.qp.pos, self.qp.rot, self.qp.vel, self.qp.ang), keep):
          t.copy_(kp)
      torch.cuda.current_stream(dev).wait_stream(side)
      g = torch.cuda.CUDAGraph()
      with torch.cuda.graph(g):
        self._launch_all()
      self._graph = g
      # the capture does not execute: the state is still the default pose

  def _launch_all(self):
    import torch
    lib, h, b = self.sys._lib, self.sys._h, self.batch
    stream = C.c_void_p(torch.cuda.current_stream(torch.device('cuda', self.sys.device)).cuda_stream)
    for io in self._ios:
      if lib.bx_step(h, b, C.byref(io), stream) != 0:
        raise RuntimeError('bx_step: ' + _lib.last_error())

  def replay(self):
    """Advances st.qp by k steps with st.acts (on torch's current stream)."""
    if self._graph is not None:
      self._graph.replay()
    else:
      self._launch_all()
    return self.qp


def load_system(name: str, device: int = 0, env=None) -> System:
  """A system from the descriptors shipped in brax_b200/systems/ (extracted
  from the reference's env configs by tools/gen_systems.py)."""
  import os
  path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'systems', name + '.json')
  return System(SystemDesc.load(path), device, env)
