"""TEST INFRASTRUCTURE ONLY: ctypes front end of oracle/brax_step_oracle.c.

The C file is the CPU restatement of the reference's `System.step`
(brax/physics/system.py:244-377).  It is the CHECKER for the CUDA path and the
timed CPU baseline; only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.  The product
(brax_b200/) never does.

  Oracle(desc, dtype)           float32 (jnp safe_norm rule) or float64
  .step(pos, rot, vel, ang, act) -> dict of outputs, batched [B, ...]
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, 'brax_step_oracle.c')

_MODE = {'pbd': 0, 'legacy_spring': 1}
_JK = {'revolute': 0, 'spherical': 1, 'spring_revolute': 2,
       'spring_universal': 3, 'spring_spherical': 4}
_AK = {'torque': 0, 'angle': 1}
_FK = {'thruster': 0, 'twister': 1}
_PK = {'capsule_plane': 0, 'capsule_capsule': 1, 'box_plane': 2, 'capsule_box': 3, 'capsule_clippedplane': 4, 'box_heightmap': 5}


def build(force: bool = False):
  """Compiles liboracle_f32.so / liboracle_f64.so next to the source.

  -ffp-contract=off keeps gcc from fusing multiply-adds, so the float32 build
  is plain IEEE single arithmetic in the reference's operation order."""
  out = []
  for name, flags in (('f32', []), ('f64', ['-DBXO_DOUBLE'])):
    lib = os.path.join(HERE, 'liboracle_%s.so' % name)
    if force or not os.path.exists(lib) or (
        os.path.getmtime(lib) < os.path.getmtime(SRC)):
      cmd = ['gcc', '-O2', '-fPIC', '-shared', '-fopenmp', '-ffp-contract=off',
             '-Wall', '-Wno-unused-function'] + flags + [SRC, '-o', lib, '-lm']
      subprocess.check_call(cmd)
    out.append(lib)
  return out


BASELINE_FLAGS = ['-O3', '-march=native', '-ffp-contract=fast', '-fno-math-errno', '-fno-trapping-math']


def build_baseline():
  """The TIMED CPU baseline build of the same port (bench.py only): `-O3 -march=native`, multiply-add
  contraction allowed -- what a performance build of this C code is.  (The checker build above stays
  `-O2 -ffp-contract=off`: plain IEEE single arithmetic in the reference's operation order.)
  -march=native binds it to the CPU it was compiled on, so it is compiled on the machine that runs
  it, keyed by the CPU's flag set.  Returns (path, flags)."""
  import hashlib
  try:
    flags_line = next(l for l in open('/proc/cpuinfo') if l.startswith('flags'))
  except Exception:
    flags_line = 'unknown'
  key = hashlib.sha1(flags_line.encode()).hexdigest()[:10]
  lib = os.path.join(HERE, 'liboracle_f32_native_%s.so' % key)
  if not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(SRC):
    cmd = ['gcc'] + BASELINE_FLAGS + ['-fPIC', '-shared', '-fopenmp', '-Wall', '-Wno-unused-function',
                                      SRC, '-o', lib, '-lm']
    subprocess.check_call(cmd)
  return lib, BASELINE_FLAGS


def _structs(R):
  class Body(C.Structure):
    _fields_ = [('mass', R), ('inv_inertia', R * 3), ('pos_mask', R * 3),
                ('rot_mask', R * 3)]

  class Joint(C.Structure):
    _fields_ = [('kind', C.c_int32), ('body_p', C.c_int32),
                ('body_c', C.c_int32), ('group', C.c_int32), ('off_p', R * 3),
                ('off_c', R * 3), ('axis_p', R * 9), ('axis_c', R * 9),
                ('limit', R * 6), ('angular_damping', R), ('scale_pos', R),
                ('scale_ang', R), ('stiffness', R), ('spring_damping', R),
                ('limit_strength', R)]

  class Actuator(C.Structure):
    _fields_ = [('kind', C.c_int32), ('joint', C.c_int32),
                ('group', C.c_int32), ('act_index', C.c_int32 * 3),
                ('strength', R)]

  class Force(C.Structure):
    _fields_ = [('kind', C.c_int32), ('body', C.c_int32),
                ('act_index', C.c_int32 * 3), ('strength', R)]

  class Pair(C.Structure):
    _fields_ = [('kind', C.c_int32), ('group', C.c_int32),
                ('two_way', C.c_int32), ('body_a', C.c_int32),
                ('body_b', C.c_int32), ('n_contacts', C.c_int32),
                ('friction', R), ('elasticity', R), ('radius_a', R),
                ('radius_b', R), ('pts', (R * 3) * 8), ('tri', ((R * 3) * 3) * 12),
                ('tri_normal', (R * 3) * 12),
                ('hm_height', C.POINTER(R)), ('hm_size', C.c_int32), ('hm_cell_size', R)]

  class Sys(C.Structure):
    _fields_ = [('mode', C.c_int32), ('substeps', C.c_int32),
                ('nb', C.c_int32), ('nj', C.c_int32), ('na', C.c_int32),
                ('nf', C.c_int32), ('np', C.c_int32), ('ngroups', C.c_int32),
                ('act_size', C.c_int32), ('safe_norm_jax', C.c_int32),
                ('dt', R), ('gravity', R * 3), ('vel_damp', R),
                ('ang_damp', R), ('erp', R), ('collide_scale', R), ('vth', R),
                ('bodies', C.POINTER(Body)), ('joints', C.POINTER(Joint)),
                ('actuators', C.POINTER(Actuator)),
                ('forces', C.POINTER(Force)), ('pairs', C.POINTER(Pair))]

  class IO(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        'pos', 'rot', 'vel', 'ang', 'act', 'pos_o', 'rot_o', 'vel_o', 'ang_o',
        'contact_vel', 'contact_ang', 'joint_vel', 'joint_ang',
        'actuator_vel', 'actuator_ang', 'contact_pos', 'contact_normal',
        'contact_penetration')]

  return Body, Joint, Actuator, Force, Pair, Sys, IO


def with_collider_friction(desc, friction: float):
  """A copy of `desc` in which EVERY collider has material.friction = `friction` -- what
  experimental/tracing/randomizers.py:26-51 (friction_randomizer) substitutes per env through
  TracedConfig: both sides of every pair (a pair's friction is friction_a * friction_b,
  colliders.py:277,315,392,459,596)."""
  import copy
  d = copy.copy(desc)
  d.pairs = [dict(p, friction_a=float(friction), friction_b=float(friction)) for p in desc.pairs]
  return d


def with_joint_offsets(desc, off_p):
  """A copy of `desc` whose joints have parent_offset (`off_p`) = off_p[j] -- what experimental/tracing/
  randomizers.py:54-97 (joint_randomizer) substitutes per env through TracedConfig."""
  import copy
  d = copy.copy(desc)
  off_p = np.asarray(off_p, np.float64).reshape(len(desc.joints), 3)
  d.joints = [dict(j, off_p=[float(x) for x in off_p[i]]) for i, j in enumerate(desc.joints)]
  return d


class FrictionOracle:
  """`Oracle.step` with per-env collider friction [B] (domain randomisation): env e is stepped by the
  oracle of `with_collider_friction(desc, friction[e])`, one Oracle per distinct value."""

  def __init__(self, desc, friction, dtype=np.float32, joint_offset=None):
    """friction: [B] or None (the descriptor's); joint_offset: [B,J,3] or None -- per-env `Joint.parent_offset`."""
    self.desc, self.dtype = desc, np.dtype(dtype)
    self.joint_offset = None if joint_offset is None else np.asarray(joint_offset).reshape(-1, len(desc.joints) * 3)
    b = len(friction) if friction is not None else self.joint_offset.shape[0]
    self.friction = None if friction is None else np.asarray(friction).reshape(-1)   # (the caller's precision: float64 values stay exact)
    # one key per env: (friction, offsets)
    self._keys = [((None if self.friction is None else float(self.friction[e])),
                   (None if self.joint_offset is None else tuple(float(x) for x in self.joint_offset[e]))) for e in range(b)]
    self._oracles = {}

  def step(self, pos, rot, vel, ang, act):
    ins = [np.asarray(x) for x in (pos, rot, vel, ang, act)]
    b = ins[0].shape[0]
    assert len(self._keys) == b
    groups = {}
    for e, k in enumerate(self._keys):
      groups.setdefault(k, []).append(e)
    out = None
    for k, idx in groups.items():
      idx = np.asarray(idx)
      o = self._oracles.get(k)
      if o is None:
        d = self.desc
        if k[0] is not None:
          d = with_collider_friction(d, k[0])
        if k[1] is not None:
          d = with_joint_offsets(d, k[1])
        o = self._oracles[k] = Oracle(d, self.dtype)
      part = o.step(*[x[idx] for x in ins])
      if out is None:
        out = {kk: np.empty((b,) + v.shape[1:], v.dtype) for kk, v in part.items()}
      for kk, v in part.items():
        out[kk][idx] = v
    return out


class Oracle:
  """CPU restatement of `System.step` for one `SystemDesc`."""

  def __init__(self, desc, dtype=np.float32, safe_norm_jax=None, baseline=False):
    """baseline=True (float32 only): the -O3 -march=native build, for TIMING the port (bench.py)."""
    f32lib, f64lib = build()
    self.dtype = np.dtype(dtype)
    is32 = self.dtype == np.float32
    R = C.c_float if is32 else C.c_double
    self.build_flags = ['-O2', '-ffp-contract=off']
    if baseline:
      if not is32:
        raise ValueError('the baseline build is float32')
      f32lib, self.build_flags = build_baseline()
    self._lib = C.CDLL(f32lib if is32 else f64lib)
    sfx = '_f32' if is32 else '_f64'
    self._step = getattr(self._lib, 'bxo_step' + sfx)
    self._step.restype = C.c_int
    self._closest = getattr(self._lib, 'bxo_closest_segment_points' + sfx)
    Body, Joint, Actuator, Force, Pair, Sys, IO = _structs(R)
    self._IO = IO
    self.desc = desc
    if safe_norm_jax is None:
      # float32 mirrors JAX (jumpy.py:184-189); float64 mirrors the numpy path
      # the golden vectors were produced on (jumpy.py:190-191).
      safe_norm_jax = is32
    cast = (lambda v: float(np.float32(v))) if is32 else float

    def arr(x):
      return [cast(v) for v in np.asarray(x, dtype=np.float64).reshape(-1)]

    nb, nj = desc.num_bodies, desc.num_joints
    bodies = (Body * max(1, nb))()
    for i, b in enumerate(desc.bodies):
      bodies[i].mass = cast(b['mass'])
      bodies[i].inv_inertia[:] = arr(b['inv_inertia'])
      bodies[i].pos_mask[:] = arr(b['pos_mask'])
      bodies[i].rot_mask[:] = arr(b['rot_mask'])
    joints = (Joint * max(1, nj))()
    for i, j in enumerate(desc.joints):
      c = joints[i]
      c.kind = _JK[j['kind']]
      c.body_p, c.body_c, c.group = j['body_p'], j['body_c'], j['group']
      c.off_p[:], c.off_c[:] = arr(j['off_p']), arr(j['off_c'])
      c.axis_p[:], c.axis_c[:] = arr(j['axis_p']), arr(j['axis_c'])
      lim = arr(j['limit'])
      c.limit[:] = lim + [0.0] * (6 - len(lim))
      for k in ('angular_damping', 'scale_pos', 'scale_ang', 'stiffness',
                'spring_damping', 'limit_strength'):
        setattr(c, k, cast(j.get(k, 0.0)))
    acts = (Actuator * max(1, len(desc.actuators)))()
    for i, a in enumerate(desc.actuators):
      acts[i].kind, acts[i].joint = _AK[a['kind']], a['joint']
      acts[i].group = a['group']
      acts[i].act_index[:] = list(a['act_index']) + [-1] * (
          3 - len(a['act_index']))
      acts[i].strength = cast(a['strength'])
    forces = (Force * max(1, len(desc.forces)))()
    for i, f in enumerate(desc.forces):
      forces[i].kind, forces[i].body = _FK[f['kind']], f['body']
      forces[i].act_index[:] = list(f['act_index'])
      forces[i].strength = cast(f['strength'])
    pairs = (Pair * max(1, len(desc.pairs)))()
    for i, p in enumerate(desc.pairs):
      c = pairs[i]
      c.kind, c.group, c.two_way = _PK[p['kind']], p['group'], int(p['two_way'])
      c.body_a, c.body_b, c.n_contacts = p['body_a'], p['body_b'], p['n_contacts']
      if is32:
        c.friction = float(np.float32(p['friction_a']) * np.float32(p['friction_b']))
        c.elasticity = float(np.float32(p['elasticity_a']) * np.float32(p['elasticity_b']))
      else:
        c.friction = p['friction_a'] * p['friction_b']
        c.elasticity = p['elasticity_a'] * p['elasticity_b']
      c.radius_a, c.radius_b = cast(p['radius_a']), cast(p['radius_b'])
      src = np.asarray(p['pts'], dtype=np.float64).reshape(-1, 3)
      for r in range(len(src)):
        c.pts[r][:] = arr(src[r])
      if p['kind'] == 'capsule_box':
        tri = np.asarray(p['tri'], dtype=np.float64).reshape(12, 3, 3)
        trn = np.asarray(p['tri_normal'], dtype=np.float64).reshape(12, 3)
        for f in range(12):
          c.tri_normal[f][:] = arr(trn[f])
          for v in range(3):
            c.tri[f][v][:] = arr(tri[f, v])
      if p['kind'] == 'box_heightmap':
        hm = np.ascontiguousarray(np.asarray(p['heightmap']['height'], dtype=np.float64).astype(self.dtype))
        self._hm_keep = getattr(self, '_hm_keep', []) + [hm]
        c.hm_height = hm.ctypes.data_as(C.POINTER(R))
        c.hm_size, c.hm_cell_size = int(hm.shape[0]), cast(p['heightmap']['cell_size'])
      if p['kind'] == 'capsule_clippedplane':
        cl = p['clip']
        c.tri[0][0][:], c.tri[0][1][:], c.tri[0][2][:] = arr(cl['normal']), arr(cl['pos']), arr(cl['x'])
        c.tri[1][0][:] = arr(cl['y'])
        c.tri[1][1][:] = arr(list(cl['halfsize']) + [0.0])
    s = Sys()
    s.mode = _MODE[desc.dynamics_mode]
    s.substeps, s.nb, s.nj = desc.substeps, nb, nj
    s.na, s.nf, s.np = len(desc.actuators), len(desc.forces), len(desc.pairs)
    s.ngroups, s.act_size = desc.num_collider_groups, desc.act_size
    s.safe_norm_jax = int(bool(safe_norm_jax))
    s.dt = cast(desc.dt_sub)
    s.gravity[:] = arr(desc.gravity)
    s.vel_damp, s.ang_damp = cast(desc.vel_damp), cast(desc.ang_damp)
    s.erp, s.collide_scale = cast(desc.baumgarte_erp), cast(desc.collide_scale)
    s.vth = cast(desc.velocity_threshold)
    s.bodies = C.cast(bodies, C.POINTER(Body))
    s.joints = C.cast(joints, C.POINTER(Joint))
    s.actuators = C.cast(acts, C.POINTER(Actuator))
    s.forces = C.cast(forces, C.POINTER(Force))
    s.pairs = C.cast(pairs, C.POINTER(Pair))
    self._sys = s
    self._keep = (bodies, joints, acts, forces, pairs)

  def angle_vel(self, rot, ang):
    """`Joint.angle_vel` of every joint (joints.py:197-226), before the free_dofs
    selection: (angle [B,nj,3], vel [B,nj,3]); slots beyond a joint's dof are 0."""
    d, dt = self.desc, self.dtype
    n, nj = d.num_bodies, len(d.joints)
    rot = np.ascontiguousarray(np.asarray(rot, dtype=dt).reshape(-1, n, 4))
    b = rot.shape[0]
    ang = np.ascontiguousarray(np.asarray(ang, dtype=dt).reshape(b, n, 3))
    angle, vel = np.zeros((b, nj, 3), dt), np.zeros((b, nj, 3), dt)
    fn = getattr(self._lib, 'bxo_angle_vel' + ('_f32' if dt == np.float32 else '_f64'))
    fn.restype = C.c_int
    rc = fn(C.byref(self._sys), C.c_int64(b), C.c_void_p(rot.ctypes.data), C.c_void_p(ang.ctypes.data),
            C.c_void_p(angle.ctypes.data), C.c_void_p(vel.ctypes.data))
    if rc != 0:
      raise RuntimeError('bxo_angle_vel failed')
    return angle, vel

  def info_contact(self, pos, rot, vel, ang):
    """`System.info(qp)` (system.py:327-340 / :378-390): info.contact (vel, ang) [B,N,3] -- the summed
    `Collider.apply(qp)` of all colliders, in both dynamics modes -- and contact_pos / contact_normal
    [B,K,3], contact_penetration [B,K]."""
    d, dt = self.desc, self.dtype
    n, k = d.num_bodies, d.num_contacts
    pos = np.ascontiguousarray(np.asarray(pos, dtype=dt).reshape(-1, n, 3))
    b = pos.shape[0]
    rot = np.ascontiguousarray(np.asarray(rot, dtype=dt).reshape(b, n, 4))
    vel = np.ascontiguousarray(np.asarray(vel, dtype=dt).reshape(b, n, 3))
    ang = np.ascontiguousarray(np.asarray(ang, dtype=dt).reshape(b, n, 3))
    out = {'contact_vel': np.zeros((b, n, 3), dt), 'contact_ang': np.zeros((b, n, 3), dt),
           'contact_pos': np.zeros((b, k, 3), dt), 'contact_normal': np.zeros((b, k, 3), dt),
           'contact_penetration': np.zeros((b, k), dt)}
    fn = getattr(self._lib, 'bxo_info_contact' + ('_f32' if dt == np.float32 else '_f64'))
    fn.restype = C.c_int
    rc = fn(C.byref(self._sys), C.c_int64(b), *[C.c_void_p(x.ctypes.data) for x in (pos, rot, vel, ang)],
            *[C.c_void_p(out[key].ctypes.data) for key in ('contact_vel', 'contact_ang', 'contact_pos',
                                                           'contact_normal', 'contact_penetration')])
    if rc != 0:
      raise RuntimeError('bxo_info_contact failed')
    return out

  def step(self, pos, rot, vel, ang, act, num_threads: int = 0):
    """One batched step.  Inputs [B,N,3|4|3|3], [B,A] (or unbatched)."""
    d, dt = self.desc, self.dtype
    n, a, k = d.num_bodies, d.act_size, d.num_contacts
    pos = np.ascontiguousarray(np.asarray(pos, dtype=dt).reshape(-1, n, 3))
    b = pos.shape[0]
    rot = np.ascontiguousarray(np.asarray(rot, dtype=dt).reshape(b, n, 4))
    vel = np.ascontiguousarray(np.asarray(vel, dtype=dt).reshape(b, n, 3))
    ang = np.ascontiguousarray(np.asarray(ang, dtype=dt).reshape(b, n, 3))
    act = np.ascontiguousarray(np.asarray(act, dtype=dt).reshape(b, a))
    out = {
        'pos': np.empty((b, n, 3), dt), 'rot': np.empty((b, n, 4), dt),
        'vel': np.empty((b, n, 3), dt), 'ang': np.empty((b, n, 3), dt),
        'contact_vel': np.empty((b, n, 3), dt),
        'contact_ang': np.empty((b, n, 3), dt),
        'joint_vel': np.empty((b, n, 3), dt),
        'joint_ang': np.empty((b, n, 3), dt),
        'actuator_vel': np.empty((b, n, 3), dt),
        'actuator_ang': np.empty((b, n, 3), dt),
        'contact_pos': np.empty((b, k, 3), dt),
        'contact_normal': np.empty((b, k, 3), dt),
        'contact_penetration': np.empty((b, k), dt),
    }
    io = self._IO()
    io.pos, io.rot = pos.ctypes.data, rot.ctypes.data
    io.vel, io.ang, io.act = vel.ctypes.data, ang.ctypes.data, act.ctypes.data
    io.pos_o, io.rot_o = out['pos'].ctypes.data, out['rot'].ctypes.data
    io.vel_o, io.ang_o = out['vel'].ctypes.data, out['ang'].ctypes.data
    for key in ('contact_vel', 'contact_ang', 'joint_vel', 'joint_ang',
                'actuator_vel', 'actuator_ang', 'contact_pos',
                'contact_normal', 'contact_penetration'):
      setattr(io, key, out[key].ctypes.data)
    rc = self._step(C.byref(self._sys), C.c_int64(b), C.byref(io),
                    C.c_int(num_threads))
    if rc != 0:
      raise RuntimeError('oracle: system exceeds compiled maxima')
    return out

  def closest_segment_points(self, a0, a1, b0, b1):
    dt = self.dtype
    v = [np.ascontiguousarray(np.asarray(x, dtype=dt)) for x in (a0, a1, b0, b1)]
    ba, bb = np.empty(3, dt), np.empty(3, dt)
    self._closest(*[C.c_void_p(x.ctypes.data) for x in v],
                  C.c_int(self._sys.safe_norm_jax),
                  C.c_void_p(ba.ctypes.data), C.c_void_p(bb.ctypes.data))
    return ba, bb

  def rotate(self, vec, quat):
    """math.rotate (math.py:25-40)."""
    dt = self.dtype
    v = np.ascontiguousarray(np.asarray(vec, dtype=dt))
    q = np.ascontiguousarray(np.asarray(quat, dtype=dt))
    out = np.empty(3, dt)
    getattr(self._lib, 'bxo_rotate' + ('_f32' if dt == np.float32 else '_f64'))(
        C.c_void_p(v.ctypes.data), C.c_void_p(q.ctypes.data), C.c_void_p(out.ctypes.data))
    return out

  def pair_contacts(self, pair_index, pos, rot, vel=None, ang=None):
    """Contact rows [C,10] = pos3 vel3 normal3 pen of one collidable pair."""
    d, dt = self.desc, self.dtype
    n = d.num_bodies
    pos = np.ascontiguousarray(np.asarray(pos, dtype=dt).reshape(n, 3))
    rot = np.ascontiguousarray(np.asarray(rot, dtype=dt).reshape(n, 4))
    vel = np.zeros((n, 3), dt) if vel is None else np.ascontiguousarray(
        np.asarray(vel, dtype=dt).reshape(n, 3))
    ang = np.zeros((n, 3), dt) if ang is None else np.ascontiguousarray(
        np.asarray(ang, dtype=dt).reshape(n, 3))
    c = d.pairs[pair_index]['n_contacts']
    out = np.empty((c, 10), dt)
    sfx = '_f32' if dt == np.float32 else '_f64'
    fn = getattr(self._lib, 'bxo_pair_contacts' + sfx)
    rc = fn(C.byref(self._sys), C.c_int(pair_index),
            *[C.c_void_p(x.ctypes.data) for x in (pos, rot, vel, ang, out)])
    if rc != 0:
      raise IndexError(pair_index)
    return out
