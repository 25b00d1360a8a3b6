"""Reads the static tables off a constructed reference `brax.System`.

`describe(sys)` walks the objects `System.__init__` built
(reference: brax/physics/system.py:53-84) and returns a `SystemDesc`.  Every
number is TAKEN from the built object rather than re-derived from the config
(SURVEY.md section 7 step 3): e.g. the collider pair list depends on quirks of
`colliders.get` (the exhausted `parents` generator at colliders.py:935, the
and/or precedence at :966-967) that a "clean" re-derivation would change.

Works with numpy-backed systems (the reference's non-jit path) and with real
JAX arrays (`np.asarray` handles both).  Anything outside the covered feature
set raises `NotImplementedError` -- fail loudly, never fall back.
"""
from typing import Any, Dict, List

import itertools

import numpy as np

from brax_b200.desc import SystemDesc

_JOINT_KINDS = {
    ('brax.physics.joints', 'Revolute'): 'revolute',
    ('brax.physics.joints', 'Spherical'): 'spherical',
    ('brax.physics.spring_joints', 'Revolute'): 'spring_revolute',
    ('brax.physics.spring_joints', 'Universal'): 'spring_universal',
    ('brax.physics.spring_joints', 'Spherical'): 'spring_spherical',
}
_PAIR_KINDS = {'capsule_plane', 'capsule_capsule', 'box_plane', 'capsule_box', 'capsule_clippedplane', 'box_heightmap'}


def _a(x) -> np.ndarray:
  return np.asarray(x, dtype=np.float64)


def _l(x) -> List[float]:
  return _a(x).reshape(-1).tolist()


def _joint_rows(joint, group: int) -> List[Dict[str, Any]]:
  key = (type(joint).__module__, type(joint).__name__)
  if key not in _JOINT_KINDS:
    raise NotImplementedError('joint type %s.%s' % key)
  kind = _JOINT_KINDS[key]
  n = len(np.asarray(joint.body_p.idx))
  rows = []
  for i in range(n):
    lim = _a(joint.limit)[i].reshape(-1, 2)
    r = {
        'kind': kind, 'group': group,
        'body_p': int(np.asarray(joint.body_p.idx)[i]),
        'body_c': int(np.asarray(joint.body_c.idx)[i]),
        'off_p': _l(_a(joint.off_p)[i]), 'off_c': _l(_a(joint.off_c)[i]),
        'axis_p': _l(_a(joint.axis_p)[i]), 'axis_c': _l(_a(joint.axis_c)[i]),
        'limit': _l(lim), 'dof': int(lim.shape[0]),
        # dofs of this joint that `Joint.angle_vel` reports (joints.py:219-225, :430-437)
        'free_dofs': int(joint.free_dofs[i]) if getattr(joint, 'free_dofs', None) is not None
                     else int(lim.shape[0]),
        'angular_damping': float(_a(joint.angular_damping)[i]),
    }
    if kind.startswith('spring_'):
      r['stiffness'] = float(_a(joint.stiffness)[i])
      r['spring_damping'] = float(_a(joint.spring_damping)[i])
      r['limit_strength'] = float(_a(joint.limit_strength)[i])
    else:
      r['scale_pos'] = float(_a(joint.scale_pos)[i])
      r['scale_ang'] = float(_a(joint.scale_ang)[i])
    rows.append(r)
  return rows


def _match_joint(aj, i: int, joints: List[Dict[str, Any]]) -> int:
  """Flat index of the joint an actuator row was gathered from
  (actuators.py:43: `jp.take(joint, [joint.index[a.joint] ...])`)."""
  bp = int(np.asarray(aj.body_p.idx)[i])
  bc = int(np.asarray(aj.body_c.idx)[i])
  key = (type(aj).__module__, type(aj).__name__)
  kind = _JOINT_KINDS[key]
  for k, j in enumerate(joints):
    if (j['kind'] == kind and j['body_p'] == bp and j['body_c'] == bc and
        np.array_equal(_a(j['off_p']), _a(aj.off_p)[i]) and
        np.array_equal(_a(j['off_c']), _a(aj.off_c)[i]) and
        np.array_equal(_a(j['axis_c']), _a(aj.axis_c)[i].reshape(-1)) and
        np.array_equal(_a(j['axis_p']), _a(aj.axis_p)[i].reshape(-1))):
      return k
  raise RuntimeError('actuator joint not found among system joints')


def _pair_rows(collider, group: int) -> List[Dict[str, Any]]:
  cull = collider.cull
  if type(cull).__name__ != 'Pairs':
    raise NotImplementedError(
        'collider cull %s (NearNeighbors is out of scope)' % type(cull).__name__)
  fn = collider.contact_fn.__name__
  if fn == 'capsule_mesh' and type(collider.cull.get()[1]).__name__ == 'TriangulatedBox':
    fn = 'capsule_box'          # ('capsule', 'box') of colliders.get (colliders.py:898-899)
  if fn not in _PAIR_KINDS:
    raise NotImplementedError('contact function %s is out of scope' % fn)
  two_way = type(collider).__name__ == 'TwoWayCollider'
  if not two_way and type(collider).__name__ != 'OneWayCollider':
    raise NotImplementedError('collider %s' % type(collider).__name__)
  ca, cb = cull.get()
  ia, ib = np.asarray(ca.body.idx), np.asarray(cb.body.idx)
  rows = []
  for i in range(len(ia)):
    r = {'kind': fn, 'group': group, 'two_way': two_way,
         'body_a': int(ia[i]), 'body_b': int(ib[i]),
         'friction_a': float(_a(ca.friction)[i]),
         'friction_b': float(_a(cb.friction)[i]),
         'elasticity_a': float(_a(ca.elasticity)[i]),
         'elasticity_b': float(_a(cb.elasticity)[i]),
         'radius_a': 0.0, 'radius_b': 0.0}
    if fn == 'capsule_plane':
      ends = _a(ca.end)[i].reshape(-1, 3)       # geometry.py:262-288
      r['n_contacts'] = int(ends.shape[0])
      r['pts'] = ends.tolist()
      r['radius_a'] = float(_a(ca.radius)[i])
    elif fn == 'capsule_clippedplane':
      # colliders.py:762-802; geometry.CapsuleEnd (:262-288) against geometry.ClippedPlane (:215-239)
      ends = _a(ca.end)[i].reshape(-1, 3)
      r['n_contacts'] = int(ends.shape[0])
      r['pts'] = ends.tolist()
      r['radius_a'] = float(_a(ca.radius)[i])
      r['clip'] = {'normal': _l(_a(cb.normal)[i]), 'pos': _l(_a(cb.pos)[i]), 'x': _l(_a(cb.x)[i]), 'y': _l(_a(cb.y)[i]),
                   'halfsize': [float(_a(cb.halfsize_x)[i]), float(_a(cb.halfsize_y)[i])]}
    elif fn == 'box_heightmap':
      # colliders.py:699-742; geometry.Box corners (:123-136) against geometry.HeightMap (:292-308)
      corners = _a(ca.corner)[i].reshape(-1, 3)
      r['n_contacts'] = int(corners.shape[0])
      r['pts'] = corners.tolist()
      r['heightmap'] = {'height': _a(cb.height)[i].tolist(), 'cell_size': float(_a(cb.cell_size)[i])}
    elif fn == 'box_plane':
      corners = _a(ca.corner)[i].reshape(-1, 3)  # geometry.py:123-136
      r['n_contacts'] = int(corners.shape[0])
      r['pts'] = corners.tolist()
    elif fn == 'capsule_box':
      # capsule_mesh against geometry.TriangulatedBox (colliders.py:822-849, geometry.py:193-199):
      # one contact row per triangle; faces / normals as the built object holds them (BaseMesh's
      # winding fix applied, collider rotation / position folded in), body-b frame
      faces = _a(cb.faces)[i].reshape(-1, 3, 3)
      r['n_contacts'] = int(faces.shape[0])
      r['pts'] = [_l(_a(ca.end)[i]), _l(_a(ca.pos)[i])]
      r['tri'] = faces.tolist()
      r['tri_normal'] = _a(cb.face_normals)[i].reshape(-1, 3).tolist()
      r['radius_a'] = float(_a(ca.radius)[i])
      if r['n_contacts'] != 12:
        raise NotImplementedError('capsule_mesh against a general mesh (only TriangulatedBox: 12 faces)')
    else:  # capsule_capsule, geometry.py:243-258
      r['n_contacts'] = 1
      r['pts'] = [_l(_a(ca.end)[i]), _l(_a(ca.pos)[i]),
                  _l(_a(cb.end)[i]), _l(_a(cb.pos)[i])]
      r['radius_a'] = float(_a(ca.radius)[i])
      r['radius_b'] = float(_a(cb.radius)[i])
    if r['n_contacts'] > 12:
      raise NotImplementedError('more than 12 contacts per pair')
    rows.append(r)
  return rows


def _reset_desc(sys) -> Dict[str, Any]:
  """The static part of `System.default_qp(joint_angle=..., joint_velocity=...)`
  (system.py:112-242) and of `default_angle` (:86-110), default_index 0:

    base     QP after the config's default qps only (:118-131)
    joints   in the depth order of :142-152: bodies, offsets, euler_to_quat of `rotation` and
             `reference_rotation` (:176-179), and `take`: the three slots of the flat joint-angle
             vector the joint reads, num_dof = the appended zero (:156-168)
    groups   trees without a default qp, lowered onto the xy plane (:211-240): per tree the
             probes of `bodies.min_z` (bodies.py:62-98) -- z = pos.z + rotate(point, rot).z -
             radius -- and `zero`: 0.0 is one of the candidates (bodies without colliders, planes)
  """
  from brax import math as bmath
  from brax.physics.base import vec_to_arr
  cfg = sys.config
  n = sys.num_bodies
  pos, rot = np.zeros((n, 3)), np.tile(np.array([1., 0., 0., 0.]), (n, 1))
  vel, ang = np.zeros((n, 3)), np.zeros((n, 3))
  default = cfg.defaults[0] if len(cfg.defaults) else None
  if default is not None:
    for dqp in default.qps:
      i = sys.body.index[dqp.name]
      pos[i], rot[i] = _a(vec_to_arr(dqp.pos)), _a(bmath.euler_to_quat(vec_to_arr(dqp.rot)))
      vel[i], ang[i] = _a(vec_to_arr(dqp.vel)), _a(vec_to_arr(dqp.ang))
  # dof slots of each joint in the flat joint-angle vector, in config order
  slots, nxt = {}, 0
  for j in cfg.joints:
    width = sum(1 for l in j.angle_limit if l.min != 0 or l.max != 0)
    slots[j.name] = (nxt, nxt + width)
    nxt += width
  num_dof = nxt
  parent_of = {j.child: j.parent for j in cfg.joints}

  def root_and_depth(body):
    hops = 0
    while body in parent_of:
      body, hops = parent_of[body], hops + 1
    return body, hops

  # parents before children: stable order by the depth of the joint's parent body
  ordered = sorted(cfg.joints, key=lambda j: root_and_depth(j.parent)[1])
  joints = []
  for j in ordered:
    b0, b1 = slots[j.name]
    take = list(range(b0, b1)) + [num_dof] * (3 - (b1 - b0))
    joints.append({
        'body_p': int(sys.body.index[j.parent]), 'body_c': int(sys.body.index[j.child]),
        'off_p': _l(vec_to_arr(j.parent_offset)), 'off_c': _l(vec_to_arr(j.child_offset)),
        'rot': _l(bmath.euler_to_quat(vec_to_arr(j.rotation))),
        'ref': _l(bmath.euler_to_quat(vec_to_arr(j.reference_rotation))),
        'take': [int(t) for t in take]})
  # free trees: bodies that are neither a joint's child nor given a default qp, with their descendants
  pinned = set(parent_of) | ({q.name for q in default.qps} if default is not None else set())
  root_idx = {b.name: [i] for i, b in enumerate(cfg.bodies) if b.name not in pinned}
  for j in ordered:
    root = root_and_depth(j.parent)[0]
    if root in root_idx:
      root_idx[root].append(int(sys.body.index[j.child]))
  groups = []
  for children in root_idx.values():
    probes, zero = [], False
    for c in children:
      body = cfg.bodies[c]
      if not body.colliders:
        zero = True
      for col in body.colliders:
        if col.HasField('sphere'):
          probes.append({'body': c, 'point': _l(vec_to_arr(col.position)), 'radius': float(col.sphere.radius)})
        elif col.HasField('capsule'):
          axis = _a(bmath.rotate(np.array([0., 0., 1.]), bmath.euler_to_quat(vec_to_arr(col.rotation))))
          length = col.capsule.length / 2 - col.capsule.radius
          for end in (-1, 1):
            probes.append({'body': c, 'point': _l(_a(vec_to_arr(col.position)) + end * axis * length),
                           'radius': float(col.capsule.radius)})
        elif col.HasField('box'):
          q = bmath.euler_to_quat(vec_to_arr(col.rotation))
          half = _a(vec_to_arr(col.box.halfsize))
          for signs in itertools.product((-1.0, 1.0), repeat=3):   # the 8 corners; the min is order-free
            corner = _a(bmath.rotate(np.array(signs) * half, q)) + _a(vec_to_arr(col.position))
            probes.append({'body': c, 'point': _l(corner), 'radius': 0.0})
        else:
          zero = True
    groups.append({'bodies': [int(c) for c in children], 'probes': probes, 'zero': zero})
  return {'num_dof': int(num_dof), 'default_angle': _l(sys.default_angle()),
          'base': {'pos': pos.tolist(), 'rot': rot.tolist(), 'vel': vel.tolist(), 'ang': ang.tolist()},
          'joints': joints, 'groups': groups}


def describe(sys, name: str = 'system', act_size: int = None) -> SystemDesc:
  """Returns the `SystemDesc` of a constructed reference `brax.System`."""
  cfg = sys.config
  # A `TracedConfig` (experimental/tracing/customize.py) is read through like the config it wraps: with a
  # scalar custom tree (one lane of the reference's vmap over the config) the built System carries the
  # substituted values and this descriptor records them; a vmapped tree (arrays) has no single descriptor
  # -- float() below fails -- and is served per env at step time instead (`System.step(friction=)`).
  mode = cfg.dynamics_mode
  integ = sys.integrator

  bodies = []
  for i in range(sys.num_bodies):
    bodies.append({
        'mass': float(_a(sys.body.mass)[i]),
        'inv_inertia': _l(_a(sys.body.inertia)[i]),
        'pos_mask': _l(_a(integ.pos_mask)[i]),
        'rot_mask': _l(_a(integ.rot_mask)[i]),
    })

  joints: List[Dict[str, Any]] = []
  for g, j in enumerate(sys.joints):
    joints.extend(_joint_rows(j, g))

  actuators = []
  for g, a in enumerate(sys.actuators):
    kind = {'Torque': 'torque', 'Angle': 'angle'}.get(type(a).__name__)
    if kind is None:
      raise NotImplementedError('actuator %s' % type(a).__name__)
    idx = np.asarray(a.act_index).reshape(len(np.asarray(a.strength)), -1)
    for i in range(idx.shape[0]):
      actuators.append({
          'kind': kind, 'group': g,
          'joint': _match_joint(a.joint, i, joints),
          'act_index': [int(v) for v in idx[i]],
          'strength': float(_a(a.strength)[i]),
      })

  forces = []
  for f in sys.forces:
    kind = {'Thruster': 'thruster', 'Twister': 'twister'}.get(type(f).__name__)
    if kind is None:
      raise NotImplementedError('force %s' % type(f).__name__)
    idx = np.asarray(f.act_index).reshape(-1, 3)
    for i in range(idx.shape[0]):
      forces.append({'kind': kind, 'body': int(np.asarray(f.body.idx)[i]),
                     'act_index': [int(v) for v in idx[i]],
                     'strength': float(_a(f.strength)[i])})

  pairs = []
  for g, c in enumerate(sys.colliders):
    pairs.extend(_pair_rows(c, g))

  if act_size is None:
    act_size = int(sys.num_joint_dof + sys.num_forces_dof)
  if sys.colliders:
    c0 = sys.colliders[0]
    erp, scale = float(c0.baumgarte_erp), float(c0.collide_scale)
    vth = float(c0.velocity_threshold)
  else:
    erp = float(cfg.baumgarte_erp * cfg.substeps / cfg.dt)
    scale = float(cfg.solver_scale_collide)
    vth = float(np.linalg.norm(_a(integ.gravity)) * float(integ.dt) * 4.0)

  qp = sys.default_qp()
  return SystemDesc(
      name=name, dynamics_mode=mode, substeps=int(cfg.substeps),
      act_size=act_size, dt=float(cfg.dt), dt_sub=float(integ.dt),
      gravity=_l(integ.gravity),
      velocity_damping=float(integ.velocity_damping),
      angular_damping=float(integ.angular_damping),
      baumgarte_erp=erp, collide_scale=scale, velocity_threshold=vth,
      num_collider_groups=len(sys.colliders),
      bodies=bodies, joints=joints, actuators=actuators, forces=forces,
      pairs=pairs, body_names=[b.name for b in cfg.bodies],
      default_qp={'pos': _a(qp.pos).tolist(), 'rot': _a(qp.rot).tolist(),
                  'vel': _a(qp.vel).tolist(), 'ang': _a(qp.ang).tolist()},
      reset=_reset_desc(sys))
