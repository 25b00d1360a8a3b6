"""Generates tests/golden/kat/: the reference's OWN known-answer tests for this path
(SURVEY.md 8(c)), restated as data (build container only).

For every selected case of /root/reference/brax/tests/{physics,geometry,colliders,
math}_test.py the script builds the system exactly as the test does (config text
taken from the imported test module, same mutations), extracts its SystemDesc
(brax_b200.extract.describe), runs the UNMODIFIED reference (numpy float64,
oracle/refsys.py) on the test's inputs and stores inputs + reference outputs.
tests/test_reference_kats.py then asserts, for the oracle (and on the GPU box for
the CUDA path), (a) the reference test's own assertion and (b) agreement with the
reference's float64 outputs.

Out of scope cases are skipped with the reason: BoxBox / BoxCapsule / Mesh /
HeightMap / ClippedPlane (hull, mesh, heightmap colliders), test_cull
(NearNeighbors), ElasticityTest (its config holds a box wall vs capsule =
capsule_mesh), test_reference_offset (default_qp forward kinematics, row N3).

usage: python oracle/gen_golden_kat.py [case-prefix ...]
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden', 'kat')


def _qp_arrays(qp):
  return {k: np.asarray(getattr(qp, k), np.float64) for k in ('pos', 'rot', 'vel', 'ang')}


def step_case(name, config, qp, act, steps=1, note=''):
  """One `sys.step` known-answer case: desc json + npz(inputs, reference outputs)."""
  import brax
  from brax_b200.extract import describe
  t0 = time.time()
  s = brax.System(config)
  act = np.asarray(act, np.float64).reshape(-1)
  a = max(1, act.shape[0])
  desc = describe(s, name, act_size=a)
  q = qp(s) if callable(qp) else qp
  rec = {('in_' + k): v for k, v in _qp_arrays(q).items()}
  rec['act'] = np.concatenate([act, np.zeros(a - act.shape[0])])
  info = None
  for _ in range(steps):
    q, info = s.step(q, act)
  rec.update({('out_' + k): v for k, v in _qp_arrays(q).items()})
  rec['steps'] = np.asarray(steps)
  # joint angles of the final state (actuator tests assert on them)
  if s.joints:
    import brax.jumpy as jp
    j0 = jp.take(s.joints[0], 0)
    _, angles = j0.axis_angle(jp.take(q, int(np.asarray(j0.body_p.idx))), jp.take(q, int(np.asarray(j0.body_c.idx))))
    rec['out_joint0_angles'] = np.asarray([float(x) for x in angles], np.float64)
  os.makedirs(OUT, exist_ok=True)
  with open(os.path.join(OUT, name + '.json'), 'w') as f:
    f.write(desc.to_json())
  np.savez_compressed(os.path.join(OUT, name + '.npz'), **rec)
  print('%-28s substeps %5d x %d steps  %.1fs %s' % (name, config.substeps, steps, time.time() - t0, note))


def physics_cases(only, module='physics_test', prefix=''):
  """module 'physics_test' (pbd) or 'physics_legacy_test' (the same classes with
  `dynamics_mode: "legacy_spring"` configs; prefix 'legacy_')."""
  brax = refsys.setup()
  from google.protobuf import text_format
  import importlib
  pt = importlib.import_module('brax.tests.' + module)
  legacy = module == 'physics_legacy_test'
  _step_case = globals()['step_case']

  def step_case(name, *args, **kw):   # noqa: F811 (prefixing wrapper)
    return _step_case(prefix + name, *args, **kw)

  def parse(text):
    return text_format.Parse(text, brax.Config())

  def want(n):
    return not only or any((prefix + n).startswith(p) for p in only)

  if want('projectile'):   # physics_test.py:32-48 / physics_legacy_test.py:32-48
    step_case('projectile', parse('''
    dt: 1 substeps: 1000
    gravity { z: -9.8 }
    bodies { name: "Ball" mass: 1 }
    defaults { qps { name: "Ball" vel {x: 1}}}
    ''' + ('dynamics_mode: "legacy_spring"' if legacy else '')), lambda s: s.default_qp(), [])
  for i, n in enumerate(('box_hits_ground', 'box_slide')):   # :67-82
    if want(n):
      step_case(n, parse(pt.BoxTest._CONFIG), lambda s, i=i: s.default_qp(i), [])
  for i, n in enumerate(('sphere_hits_ground', 'sphere_roll')):   # :275-287
    if want(n):
      step_case(n, parse(pt.SphereTest._CONFIG), lambda s, i=i: s.default_qp(i), [])
  if want('capsule_hits_ground'):   # :330-338
    step_case('capsule_hits_ground', parse(pt.CapsuleTest._CONFIG), lambda s: s.default_qp(0), [])
  if want('capsule_hits_capsule'):   # :340-349
    c = parse(pt.CapsuleTest._CONFIG)
    c.dt, c.substeps = 2.0, (1000 if legacy else 400)   # physics_legacy_test.py:298-299
    step_case('capsule_hits_capsule', c, lambda s: s.default_qp(1), [])
  for k, (mass, radius, vel) in enumerate(((2.0, 0.125, 0.0625), (5.0, 0.125, 0.03125), (1.0, 0.0625, 0.1))):
    n = 'pendulum_%d' % k    # :497-523
    if want(n):
      c = parse(pt.JointTest._CONFIG)
      c.dt = float(2 * np.pi * np.sqrt((.4 * radius**2 + 1.) / 9.8))
      c.bodies[1].mass = mass
      for ax in 'xyz':
        setattr(c.bodies[1].inertia, ax, .4 * mass * radius**2)
      qp = brax.QP(pos=np.array([[0., 0., 0.], [0., 0., -1.]]), rot=np.array([[1., 0., 0., 0.], [1., 0., 0., 0.]]),
                   vel=np.array([[0., 0., 0.], [0., vel, 0.]]), ang=np.array([[0., 0., 0.], [vel, 0., 0.]]))
      step_case(n, c, qp, [])
  for t in (15., 30., 45., 90.):   # :624-636
    n = 'actuator1d_%d' % int(t)
    if want(n):
      step_case(n, parse(pt.Actuator1DTest._CONFIG), lambda s: s.default_qp(), [t])
  for t1, t2 in ((15., 30.), (-45., 80.), (120., -60.), (-35., -52.)):   # :667-679
    n = 'actuator2d_%d_%d' % (int(t1), int(t2))
    if want(n):
      step_case(n, parse(pt.Actuator2DTest._CONFIG), lambda s: s.default_qp(), [t1, t2])
  for limits, tq in (((35, 40, 75), (1, 1, 1)), ((80, 45, 30), (1, 0, 0))):   # :722-742 (2 of the 12 combinations:
    n = 'actuator3d_%d_%d_%d_t%d%d%d' % (limits + tq)                           # 8000 substeps each on numpy)
    if want(n):
      c = parse(pt.Actuator3DTest._CONFIG)
      for al, lim in zip(c.joints[0].angle_limit, limits):
        al.min, al.max = -lim, lim
      step_case(n, c, lambda s: s.default_qp(), list(tq))
  for f in (() if legacy else (1, 5, 10)):   # :813-831 (physics_legacy_test.py has no ForceTest)
    if want('thruster_%d' % f):
      step_case('thruster_%d' % f, parse(pt.ForceTest._CONFIG), lambda s: s.default_qp(),
                f * np.array([1., 0., 0., 0., 0., 0.]))
    if want('twister_%d' % f):
      step_case('twister_%d' % f, parse(pt.ForceTest._CONFIG), lambda s: s.default_qp(),
                f * np.array([0., 0., 0., 1., 0., 0.]))


def unit_cases():
  """geometry_test.py:214-311 (closest segment points: 7 named + 100 random cases),
  colliders_test.py:54-135 (capsule_capsule), math_test.py:25-37 (rotate)."""
  brax = refsys.setup()
  from google.protobuf import text_format
  import importlib
  from brax.physics import geometry
  from brax import math as bmath
  import brax.jumpy as jp
  from brax_b200.extract import describe
  gt = importlib.import_module('brax.tests.geometry_test')
  ct = importlib.import_module('brax.tests.colliders_test')
  A = np.array
  named = [
      (A([0.73432405, 0.12372768, 0.20272314]), A([1.10600128, 0.88555209, 0.65209485]),
       A([0.85599262, 0.61736299, 0.9843583]), A([1.84270939, 0.92891793, 1.36343326])),
      (A([0., 0., -1.]), A([0., 0., 1.]), A([-1., 0., 0.]), A([1., 0., 0.])),
      (A([0.2, 0.2, 0.]), A([1., 1., 0.]), A([0.2, 0.4, 0.]), A([1., 2., 0.])),
      (A([0., 0., -1.]), A([0., 0., 1.]), A([1., 0., -1.]), A([1., 0., 1.])),
      (A([0., 0., -1.]), A([0., 0., 1.]), A([1., 0., 1.]), A([1., 0., 3.])),
      (A([0., 0., -1.]), A([0., 0., -1.]), A([1., 0., 0.1]), A([1., 0., 0.1])),
      (A([0., 0., -1.]), A([0., 0., 1.]), A([0., 0., -1.]), A([0., 0., 1.])),
  ]
  segs = list(named)
  for i in range(10):
    for j in range(10):
      a0, a1 = gt._get_rand_line_segment(i)
      b0, b1 = gt._get_rand_line_segment(j)
      segs.append((A(a0, float), A(a1, float), A(b0, float), A(b1, float)))
  seg_in = np.stack([np.stack(s) for s in segs])
  seg_out = np.stack([np.stack([np.asarray(x, np.float64) for x in
                                geometry.closest_segment_to_segment_points(*s)]) for s in segs])
  rec = {'seg_in': seg_in, 'seg_out': seg_out}
  # capsule_capsule
  for tag, text, zero in (('cc1', ct.CapsuleCapsuleColliderFnTest._CONFIG, False),
                          ('cc2', ct.CapsuleCapsuleColliderFnTest._CONFIG2, False),
                          ('cc2zero', ct.CapsuleCapsuleColliderFnTest._CONFIG2, True)):
    config = text_format.Parse(text, brax.Config())
    s = brax.System(config)
    capsules = geometry.Capsule(config.bodies, s.body)
    c1, c2 = jp.take(capsules, 0), jp.take(capsules, 1)
    if zero:
      c1.end = np.array([0., 0., 0.])
      c2.end = np.array([0., 0., 0.])
    qp = s.default_qp()
    contact = brax.physics.colliders.capsule_capsule(c1, c2, jp.take(qp, 0), jp.take(qp, 1))
    rec[tag + '_pos'] = np.asarray(qp.pos, np.float64)
    rec[tag + '_rot'] = np.asarray(qp.rot, np.float64)
    rec[tag + '_contact'] = np.concatenate([np.asarray(contact.pos, np.float64).reshape(-1),
                                            np.asarray(contact.vel, np.float64).reshape(-1),
                                            np.asarray(contact.normal, np.float64).reshape(-1),
                                            np.asarray(contact.penetration, np.float64).reshape(-1)])
    if not zero:
      with open(os.path.join(OUT, 'unit_' + tag + '.json'), 'w') as f:
        f.write(describe(s, 'unit_' + tag, act_size=1).to_json())
  # rotate
  quat = np.asarray(bmath.euler_to_quat(np.array([0., 45., 0.])), np.float64)
  rec['rot_vec'] = np.array([[0., 0., 1.], [1., 1., 1.]])
  rec['rot_quat'] = np.stack([quat, np.array([1., 0., 0., 0.])])
  rec['rot_out'] = np.stack([np.asarray(bmath.rotate(rec['rot_vec'][i], rec['rot_quat'][i]), np.float64)
                             for i in range(2)])
  os.makedirs(OUT, exist_ok=True)
  np.savez_compressed(os.path.join(OUT, 'unit.npz'), **rec)
  print('unit cases: %d segment pairs, 3 capsule_capsule, 2 rotate' % len(segs))


if __name__ == '__main__':
  only = sys.argv[1:]
  if not only or 'unit' in only:
    unit_cases()
  sel = [o for o in only if o != 'unit'] if only != ['unit'] else ['__none__']
  if not sel or any(not o.startswith('legacy_') for o in sel):
    physics_cases([o for o in sel if not o.startswith('legacy_')] or (['__none__'] if sel else []))
  if not sel or any(o.startswith('legacy_') for o in sel):
    # physics_legacy_test.py: the same known-answer tests in dynamics_mode legacy_spring
    physics_cases([o for o in sel if o.startswith('legacy_')] or (['__none__'] if sel else []),
                  'physics_legacy_test', 'legacy_')
