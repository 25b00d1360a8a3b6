"""Static system description: the host-side mirror of `bx_system_desc`.

A `SystemDesc` holds every number `brax.System.__init__` derives from a config
(reference: brax/physics/system.py:53-84) as plain Python/numpy data:

  * it is what `brax_b200.extract.describe(sys)` reads off a constructed
    reference `System`,
  * it round-trips through JSON (values stored as float64 so the float64
    oracle sees un-rounded tables),
  * `to_c()` packs it into the ctypes structs declared in
    include/brax_b200.h (float32 -- what the reference computes in under JAX
    defaults).

Nothing here does physics.
"""
import ctypes as C
import dataclasses
import json
from typing import Any, Dict, List

import numpy as np

ABI_VERSION = 14

MODE = {'pbd': 0, 'legacy_spring': 1}
JOINT_KIND = {'revolute': 0, 'spherical': 1, 'spring_revolute': 2,
              'spring_universal': 3, 'spring_spherical': 4}
ACT_KIND = {'torque': 0, 'angle': 1}
FORCE_KIND = {'thruster': 0, 'twister': 1}
PAIR_KIND = {'capsule_plane': 0, 'capsule_capsule': 1, 'box_plane': 2, 'capsule_box': 3, 'capsule_clippedplane': 4,
             'box_heightmap': 5}

MAX_BODIES, MAX_JOINTS, MAX_ACTUATORS = 64, 64, 64
MAX_FORCES, MAX_PAIRS, MAX_CONTACTS, MAX_GROUPS = 32, 64, 128, 8
MAX_PROBES, MAX_TREES = 128, 8


class CBody(C.Structure):
  _fields_ = [('mass', C.c_float), ('inv_inertia', C.c_float * 3),
              ('pos_mask', C.c_float * 3), ('rot_mask', C.c_float * 3)]


class CJoint(C.Structure):
  _fields_ = [('kind', C.c_int32), ('body_p', C.c_int32),
              ('body_c', C.c_int32), ('group', C.c_int32),
              ('off_p', C.c_float * 3), ('off_c', C.c_float * 3),
              ('axis_p', C.c_float * 9), ('axis_c', C.c_float * 9),
              ('limit', C.c_float * 6), ('angular_damping', C.c_float),
              ('scale_pos', C.c_float), ('scale_ang', C.c_float),
              ('stiffness', C.c_float), ('spring_damping', C.c_float),
              ('limit_strength', C.c_float)]


class CActuator(C.Structure):
  _fields_ = [('kind', C.c_int32), ('joint', C.c_int32), ('group', C.c_int32),
              ('act_index', C.c_int32 * 3), ('strength', C.c_float)]


class CForce(C.Structure):
  _fields_ = [('kind', C.c_int32), ('body', C.c_int32),
              ('act_index', C.c_int32 * 3), ('strength', C.c_float)]


class CPair(C.Structure):
  _fields_ = [('kind', C.c_int32), ('group', C.c_int32),
              ('two_way', C.c_int32), ('body_a', C.c_int32),
              ('body_b', C.c_int32), ('n_contacts', C.c_int32),
              ('friction', C.c_float), ('elasticity', C.c_float),
              ('radius_a', C.c_float), ('radius_b', C.c_float),
              ('pts', (C.c_float * 3) * 8), ('tri', ((C.c_float * 3) * 3) * 12),
              ('tri_normal', (C.c_float * 3) * 12),
              ('hm_height', C.POINTER(C.c_float)), ('hm_size', C.c_int32), ('hm_cell_size', C.c_float)]


class CResetJoint(C.Structure):
  _fields_ = [('body_p', C.c_int32), ('body_c', C.c_int32), ('take', C.c_int32 * 3),
              ('off_p', C.c_float * 3), ('off_c', C.c_float * 3), ('rot', C.c_float * 4), ('ref', C.c_float * 4)]


class CResetProbe(C.Structure):
  _fields_ = [('body', C.c_int32), ('tree', C.c_int32), ('point', C.c_float * 3), ('radius', C.c_float)]


class CResetDesc(C.Structure):
  _fields_ = [('struct_size', C.c_uint32), ('num_dof', C.c_int32), ('num_joints', C.c_int32),
              ('num_probes', C.c_int32), ('num_trees', C.c_int32),
              ('base_pos', (C.c_float * 3) * MAX_BODIES), ('base_rot', (C.c_float * 4) * MAX_BODIES),
              ('base_vel', (C.c_float * 3) * MAX_BODIES), ('base_ang', (C.c_float * 3) * MAX_BODIES),
              ('joints', CResetJoint * MAX_JOINTS), ('probes', CResetProbe * MAX_PROBES),
              ('body_tree', C.c_int32 * MAX_BODIES), ('tree_zero', C.c_int32 * MAX_TREES)]


class CSystemDesc(C.Structure):
  _fields_ = [('struct_size', C.c_uint32), ('abi_version', C.c_uint32),
              ('dynamics_mode', C.c_int32), ('substeps', C.c_int32),
              ('num_bodies', C.c_int32), ('num_joints', C.c_int32),
              ('num_actuators', C.c_int32), ('num_forces', C.c_int32),
              ('num_pairs', C.c_int32), ('num_collider_groups', C.c_int32),
              ('act_size', C.c_int32), ('dt_sub', C.c_float),
              ('gravity', C.c_float * 3), ('vel_damp', C.c_float),
              ('ang_damp', C.c_float), ('baumgarte_erp', C.c_float),
              ('collide_scale', C.c_float), ('velocity_threshold', C.c_float),
              ('bodies', C.POINTER(CBody)), ('joints', C.POINTER(CJoint)),
              ('actuators', C.POINTER(CActuator)),
              ('forces', C.POINTER(CForce)), ('pairs', C.POINTER(CPair)),
              ('env', C.c_void_p), ('reset', C.c_void_p)]


class CStepIO(C.Structure):
  _fields_ = [(n, C.c_void_p) for n in (
      'pos', 'rot', 'vel', 'ang', 'act', 'pos_o', 'rot_o', 'vel_o', 'ang_o',
      'contact_vel', 'contact_ang', 'joint_vel', 'joint_ang', 'actuator_vel',
      'actuator_ang', 'contact_pos', 'contact_normal', 'contact_penetration',
      'joint_angle', 'joint_velocity', 'obs', 'reward', 'done', 'metrics', 'env_rand',
      'first_pos', 'first_rot', 'first_vel', 'first_ang', 'first_obs', 'done_in', 'steps_in',
      'steps_out', 'truncation', 'policy', 'obs0', 'act_out', 'friction', 'joint_offset')] + [('episode_length', C.c_float), ('observe', C.c_int32), ('action_repeat', C.c_int32)]


POLICY_MAX_LAYERS = 8
POLICY_ACT = {'relu': 0, 'swish': 1, 'tanh': 2}


class CPolicy(C.Structure):
  _fields_ = [('struct_size', C.c_uint32), ('num_layers', C.c_int32), ('sizes', C.c_int32 * (POLICY_MAX_LAYERS + 1)),
              ('activation', C.c_int32), ('min_std', C.c_float), ('params', C.c_void_p), ('obs_mean', C.c_void_p),
              ('obs_std', C.c_void_p), ('noise', C.c_void_p)]


def _f(x):
  return [float(v) for v in np.asarray(x, dtype=np.float64).reshape(-1)]


@dataclasses.dataclass
class SystemDesc:
  """Plain-data description of one brax v1 system (see module docstring)."""
  name: str
  dynamics_mode: str
  substeps: int
  act_size: int
  dt: float                 # config.dt
  dt_sub: float             # config.dt / substeps   (integrators.py:45)
  gravity: List[float]
  velocity_damping: float
  angular_damping: float
  baumgarte_erp: float      # already scaled by substeps/dt (colliders.py:109)
  collide_scale: float
  velocity_threshold: float
  num_collider_groups: int
  bodies: List[Dict[str, Any]]
  joints: List[Dict[str, Any]]
  actuators: List[Dict[str, Any]]
  forces: List[Dict[str, Any]]
  pairs: List[Dict[str, Any]]
  body_names: List[str] = dataclasses.field(default_factory=list)
  # convenience payload for tests / bench (not consumed by the kernel):
  default_qp: Dict[str, Any] = dataclasses.field(default_factory=dict)
  reset: Dict[str, Any] = dataclasses.field(default_factory=dict)

  # ---- derived sizes -------------------------------------------------------
  @property
  def num_bodies(self) -> int:
    return len(self.bodies)

  @property
  def num_joints(self) -> int:
    return len(self.joints)

  @property
  def num_contacts(self) -> int:
    return sum(p['n_contacts'] for p in self.pairs)

  @property
  def vel_damp(self) -> float:
    # exp() is evaluated on the host in float64 at trace time and then rounded
    # (integrators.py:89,92).
    return float(np.exp(self.velocity_damping * self.dt_sub))

  @property
  def ang_damp(self) -> float:
    return float(np.exp(self.angular_damping * self.dt_sub))

  def bytes_per_env_step(self) -> int:
    """Algorithmic HBM bytes of one env-step:
    4 * [(13N + A) in + 13N out + 18N info + 7K contacts]."""
    n, a, k = self.num_bodies, self.act_size, self.num_contacts
    return 4 * ((13 * n + a) + 13 * n + 18 * n + 7 * k)

  # ---- JSON ----------------------------------------------------------------
  def to_json(self) -> str:
    return json.dumps(dataclasses.asdict(self), indent=1)

  @classmethod
  def from_json(cls, text: str) -> 'SystemDesc':
    return cls(**json.loads(text))

  @classmethod
  def load(cls, path: str) -> 'SystemDesc':
    with open(path) as f:
      return cls.from_json(f.read())

  def validate(self):
    """Raises ValueError for anything the kernels do not cover."""
    if self.dynamics_mode not in MODE:
      raise ValueError('unknown dynamics_mode %r' % self.dynamics_mode)
    lim = [('bodies', self.num_bodies, MAX_BODIES),
           ('joints', self.num_joints, MAX_JOINTS),
           ('actuators', len(self.actuators), MAX_ACTUATORS),
           ('forces', len(self.forces), MAX_FORCES),
           ('pairs', len(self.pairs), MAX_PAIRS),
           ('contacts', self.num_contacts, MAX_CONTACTS),
           ('collider groups', self.num_collider_groups, MAX_GROUPS)]
    for what, n, m in lim:
      if n > m:
        raise ValueError('%d %s exceeds the compiled maximum %d' % (n, what, m))
    for j in self.joints:
      if j['kind'] not in JOINT_KIND:
        raise ValueError('unsupported joint kind %r' % j['kind'])
    for p in self.pairs:
      if p['kind'] not in PAIR_KIND:
        raise ValueError('unsupported collider pair %r' % p['kind'])

  # ---- C packing -----------------------------------------------------------
  def to_c(self, env=None):
    """Returns (CSystemDesc, keepalive) with float32-rounded tables.  env: optional
    `brax_b200.envs.EnvDesc` (the env epilogue, bx_env_desc)."""
    self.validate()
    bodies = (CBody * max(1, self.num_bodies))()
    for i, b in enumerate(self.bodies):
      bodies[i].mass = b['mass']
      bodies[i].inv_inertia[:] = _f(b['inv_inertia'])
      bodies[i].pos_mask[:] = _f(b['pos_mask'])
      bodies[i].rot_mask[:] = _f(b['rot_mask'])
    joints = (CJoint * max(1, self.num_joints))()
    for i, j in enumerate(self.joints):
      c = joints[i]
      c.kind = JOINT_KIND[j['kind']]
      c.body_p, c.body_c, c.group = j['body_p'], j['body_c'], j['group']
      c.off_p[:] = _f(j['off_p'])
      c.off_c[:] = _f(j['off_c'])
      c.axis_p[:] = _f(j['axis_p'])
      c.axis_c[:] = _f(j['axis_c'])
      lim = _f(j['limit'])
      c.limit[:] = lim + [0.0] * (6 - len(lim))
      for k in ('angular_damping', 'scale_pos', 'scale_ang', 'stiffness',
                'spring_damping', 'limit_strength'):
        setattr(c, k, float(j.get(k, 0.0)))
    acts = (CActuator * max(1, len(self.actuators)))()
    for i, a in enumerate(self.actuators):
      acts[i].kind = ACT_KIND[a['kind']]
      acts[i].joint, acts[i].group = a['joint'], a['group']
      idx = list(a['act_index']) + [-1] * (3 - len(a['act_index']))
      acts[i].act_index[:] = idx
      acts[i].strength = a['strength']
    forces = (CForce * max(1, len(self.forces)))()
    for i, f in enumerate(self.forces):
      forces[i].kind = FORCE_KIND[f['kind']]
      forces[i].body = f['body']
      forces[i].act_index[:] = list(f['act_index'])
      forces[i].strength = f['strength']
    pairs = (CPair * max(1, len(self.pairs)))()
    hm_keep = []
    for i, p in enumerate(self.pairs):
      c = pairs[i]
      c.kind = PAIR_KIND[p['kind']]
      c.group, c.two_way = p['group'], int(p['two_way'])
      c.body_a, c.body_b, c.n_contacts = p['body_a'], p['body_b'], p['n_contacts']
      # the reference multiplies the two float32 material values at run time
      # (colliders.py:317,393); do the same product in float32 here.
      c.friction = float(np.float32(p['friction_a']) * np.float32(p['friction_b']))
      c.elasticity = float(
          np.float32(p['elasticity_a']) * np.float32(p['elasticity_b']))
      c.radius_a, c.radius_b = p['radius_a'], p['radius_b']
      pts = np.zeros((8, 3))
      src = np.asarray(p['pts'], dtype=np.float64).reshape(-1, 3)
      pts[:len(src)] = src
      for r in range(8):
        c.pts[r][:] = _f(pts[r])
      if p['kind'] == 'capsule_box':
        tri = np.asarray(p['tri'], dtype=np.float64).reshape(12, 3, 3)
        trn = np.asarray(p['tri_normal'], dtype=np.float64).reshape(12, 3)
        for f in range(12):
          c.tri_normal[f][:] = _f(trn[f])
          for v in range(3):
            c.tri[f][v][:] = _f(tri[f, v])
      if p['kind'] == 'box_heightmap':
        hm = np.ascontiguousarray(np.asarray(p['heightmap']['height'], dtype=np.float32))
        assert hm.ndim == 2 and hm.shape[0] == hm.shape[1]
        hm_keep.append(hm)
        c.hm_height = hm.ctypes.data_as(C.POINTER(C.c_float))
        c.hm_size, c.hm_cell_size = int(hm.shape[0]), float(p['heightmap']['cell_size'])
      if p['kind'] == 'capsule_clippedplane':
        cl = p['clip']      # geometry.ClippedPlane: normal, pos, x, y, halfsize (body-b frame)
        c.tri[0][0][:], c.tri[0][1][:], c.tri[0][2][:] = _f(cl['normal']), _f(cl['pos']), _f(cl['x'])
        c.tri[1][0][:] = _f(cl['y'])
        c.tri[1][1][:] = _f(list(cl['halfsize']) + [0.0])
    d = CSystemDesc()
    d.struct_size = C.sizeof(CSystemDesc)
    d.abi_version = ABI_VERSION
    d.dynamics_mode = MODE[self.dynamics_mode]
    d.substeps = self.substeps
    d.num_bodies, d.num_joints = self.num_bodies, self.num_joints
    d.num_actuators, d.num_forces = len(self.actuators), len(self.forces)
    d.num_pairs = len(self.pairs)
    d.num_collider_groups = self.num_collider_groups
    d.act_size = self.act_size
    d.dt_sub = self.dt_sub
    d.gravity[:] = _f(self.gravity)
    d.vel_damp, d.ang_damp = self.vel_damp, self.ang_damp
    d.baumgarte_erp = self.baumgarte_erp
    d.collide_scale = self.collide_scale
    d.velocity_threshold = self.velocity_threshold
    d.bodies = C.cast(bodies, C.POINTER(CBody))
    d.joints = C.cast(joints, C.POINTER(CJoint))
    d.actuators = C.cast(acts, C.POINTER(CActuator))
    d.forces = C.cast(forces, C.POINTER(CForce))
    d.pairs = C.cast(pairs, C.POINTER(CPair))
    cenv = env.to_c(self) if env is not None else None
    d.env = C.cast(C.pointer(cenv), C.c_void_p) if cenv is not None else None
    creset = self.reset_to_c()
    d.reset = C.cast(C.pointer(creset), C.c_void_p) if creset is not None else None
    return d, (bodies, joints, acts, forces, pairs, cenv, creset, hm_keep)

  def reset_to_c(self):
    """bx_reset_desc of `self.reset` (extract._reset_desc), or None when the descriptor has none."""
    r = self.reset
    if not r:
      return None
    probes = [(p, t) for t, g in enumerate(r['groups']) for p in g['probes']]
    if len(r['joints']) > MAX_JOINTS or len(probes) > MAX_PROBES or len(r['groups']) > MAX_TREES:
      raise ValueError('reset tables exceed the compiled maxima')
    c = CResetDesc()
    c.struct_size = C.sizeof(CResetDesc)
    c.num_dof, c.num_joints = int(r['num_dof']), len(r['joints'])
    c.num_probes, c.num_trees = len(probes), len(r['groups'])
    for k, w in (('pos', 3), ('rot', 4), ('vel', 3), ('ang', 3)):
      arr = getattr(c, 'base_' + k)
      for b in range(self.num_bodies):
        arr[b][:] = _f(r['base'][k][b])
    for i, j in enumerate(r['joints']):
      cj = c.joints[i]
      cj.body_p, cj.body_c = j['body_p'], j['body_c']
      cj.take[:] = [int(t) for t in j['take']]
      cj.off_p[:], cj.off_c[:] = _f(j['off_p']), _f(j['off_c'])
      cj.rot[:], cj.ref[:] = _f(j['rot']), _f(j['ref'])
    for i, (p, t) in enumerate(probes):
      c.probes[i].body, c.probes[i].tree = p['body'], t
      c.probes[i].point[:] = _f(p['point'])
      c.probes[i].radius = p['radius']
    c.body_tree[:] = [-1] * MAX_BODIES
    for t, g in enumerate(r['groups']):
      c.tree_zero[t] = int(bool(g['zero']))
      for b in g['bodies']:
        c.body_tree[b] = t
    return c
