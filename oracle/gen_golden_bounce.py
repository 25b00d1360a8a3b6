"""Generates tests/golden/bounce/: reference trajectories that exercise RESTITUTION
(elasticity > 0) -- the branch no env config of the reference reaches (build container only).

Source: `ElasticityTest` (/root/reference/brax/tests/physics_test.py:835-914).  Its config is
used as is except that the far-away `boxwall` body (a box vs capsule pair = capsule_mesh,
out of scope; 10 m from everything in the vertical tests) is dropped:

  ball_ball_<mode>_e<E>_<free|frozen>   test_ball_bounce_vertical(_frozen) (:880-914):
        default_qp(1), dt = (2*5/9.8)/100, elasticity E in {0, .5, 1}; a capsule-capsule
        pair (TwoWayCollider; OneWayCollider when the lower ball is frozen) + two
        capsule-plane pairs; 400 (free) / 110 (frozen) steps
  ball_plane_<mode>_e50                 a ball thrown onto the ground plane, elasticity .5
        (the one-way capsule_plane restitution branch, colliders.py:411-430)

in dynamics modes pbd and legacy_spring.  The UNMODIFIED reference (numpy float64,
oracle/refsys.py) is rolled forward and EVERY step's input state and outputs (QP + Info)
are stored, so the tests check single steps from the reference's own states (tight) as
well as the reference test's assertion on the chained result.

usage: python oracle/gen_golden_bounce.py [name-prefix ...]
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import refsys  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden', 'bounce')
INFO = ('contact_vel', 'contact_ang', 'joint_vel', 'joint_ang', 'actuator_vel', 'actuator_ang',
        'contact_pos', 'contact_normal', 'contact_penetration')


def _info_arrays(info):
  f = lambda x: np.asarray(x, np.float64)  # noqa: E731
  return {'contact_vel': f(info.contact.vel), 'contact_ang': f(info.contact.ang),
          'joint_vel': f(info.joint.vel), 'joint_ang': f(info.joint.ang),
          'actuator_vel': f(info.actuator.vel), 'actuator_ang': f(info.actuator.ang),
          'contact_pos': f(info.contact_pos), 'contact_normal': f(info.contact_normal),
          'contact_penetration': f(info.contact_penetration)}


def roll(name, config, qp_index, steps, note=''):
  import brax
  from brax_b200.extract import describe
  t0 = time.time()
  s = brax.System(config)
  desc = describe(s, name, act_size=1)
  qp = s.default_qp(qp_index)
  rec = {k: [] for k in ('pos', 'rot', 'vel', 'ang')}
  rec.update({'out_' + k: [] for k in ('pos', 'rot', 'vel', 'ang') + INFO})
  act = np.zeros(0)
  for _ in range(steps):
    for k in ('pos', 'rot', 'vel', 'ang'):
      rec[k].append(np.asarray(getattr(qp, k), np.float64))
    qp, info = s.step(qp, act)
    for k in ('pos', 'rot', 'vel', 'ang'):
      rec['out_' + k].append(np.asarray(getattr(qp, k), np.float64))
    for k, v in _info_arrays(info).items():
      rec['out_' + k].append(v)
  rec = {k: np.stack(v) for k, v in rec.items()}
  rec['act'] = np.zeros((steps, 1))
  rec['elasticity'] = np.asarray(float(config.elasticity))
  os.makedirs(OUT, exist_ok=True)
  with open(os.path.join(OUT, name + '.json'), 'w') as f:
    f.write(desc.to_json())
  np.savez_compressed(os.path.join(OUT, name + '.npz'), **rec)
  vz = rec['out_vel'][:, 0, 2]
  print('%-34s %4d steps  %.1fs  final vel.z %+.4f  max vel.z %+.4f %s' % (
      name, steps, time.time() - t0, vz[-1], vz[5:].max(), note))


def main(only):
  brax = refsys.setup()
  from google.protobuf import text_format
  import importlib
  pt = importlib.import_module('brax.tests.physics_test')

  def want(n):
    return not only or any(n.startswith(p) for p in only)

  def elasticity_config(mode):
    c = text_format.Parse(pt.ElasticityTest._CONFIG, brax.Config())
    assert c.bodies[2].name == 'boxwall'
    del c.bodies[2]
    for d in c.defaults:
      keep = [q for q in d.qps if q.name != 'boxwall']
      del d.qps[:]
      d.qps.extend(keep)
    if mode == 'spring':
      c.dynamics_mode = 'legacy_spring'
    return c

  for mode in ('pbd', 'spring'):
    for e in (0.0, 0.5, 1.0):
      for frozen in (False, True):
        n = 'ball_ball_%s_e%d_%s' % (mode, int(100 * e), 'frozen' if frozen else 'free')
        if not want(n):
          continue
        c = elasticity_config(mode)
        c.elasticity = e
        if frozen:
          c.bodies[1].frozen.all = True      # physics_test.py:901
        c.dt = (2 * 5 / 9.8) / 100.          # :886-887
        roll(n, c, 1, 110 if frozen else 400)
    # ElasticityTest.test_ball_bounce (physics_test.py:863-878): the ball flies into the frozen box wall
    # -- capsule_mesh against geometry.TriangulatedBox (colliders.py:822-849), one-way -- with the
    # config AS IS (boxwall included), default_qp(0), 100 steps
    for e in (0.0, 0.5, 1.0):
      n = 'ball_wall_%s_e%d' % (mode, int(100 * e))
      if want(n):
        c = text_format.Parse(pt.ElasticityTest._CONFIG, brax.Config())
        c.elasticity = e
        if mode == 'spring':
          c.dynamics_mode = 'legacy_spring'
        roll(n, c, 0, 100)
    # a tilted capsule dropped onto a frozen box (edge / face contacts of the triangulated box, one-way)
    # and a ball dropped onto a FREE box resting on the plane (two-way capsule_mesh + box_plane)
    n = 'capsule_on_box_%s' % mode
    if want(n):
      c = text_format.Parse('''
        dt: 0.02 substeps: 8 friction: 0.6 elasticity: 0.3
        gravity { z: -9.8 }
        bodies {
          name: "cap" mass: 1
          colliders { rotation { y: 60 } capsule { radius: .15 length: 1.0 } }
          inertia { x: 0.05 y: 0.05 z: 0.02 }
        }
        bodies {
          name: "ball" mass: 0.5
          colliders { capsule { radius: .25 length: 0.5 } }
          inertia { x: 0.02 y: 0.02 z: 0.02 }
        }
        bodies {
          name: "block" mass: 1
          colliders { rotation { z: 20 } box { halfsize { x: 0.6 y: 0.4 z: 0.3 } } }
          inertia { x: 1 y: 1 z: 1 }
          frozen { all: true }
        }
        bodies {
          name: "crate" mass: 2
          colliders { box { halfsize { x: 0.4 y: 0.4 z: 0.4 } } }
          inertia { x: 0.3 y: 0.3 z: 0.3 }
        }
        bodies { name: "Ground" frozen: { all: true } colliders { plane {}}}
        collide_include { first: "cap" second: "block" }
        collide_include { first: "cap" second: "Ground" }
        collide_include { first: "ball" second: "crate" }
        collide_include { first: "ball" second: "Ground" }
        collide_include { first: "crate" second: "Ground" }
        defaults {
          qps { name: "cap" pos { x: 0.3 y: 0.1 z: 1.2 } vel { z: -1.0 } ang { x: 0.5 } }
          qps { name: "ball" pos { x: 3.1 y: 0.05 z: 1.6 } vel { z: -2.0 } }
          qps { name: "block" pos { z: 0.3 } }
          qps { name: "crate" pos { x: 3.0 z: 0.4 } }
        }
      ''', brax.Config())
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 80)
    # CapsuleClippedPlaneTest (physics_test.py:426-476): three spheres, one lands on the clipped plane, two
    # miss it and land on the ground -- the config AS IS except dt: the reference takes ONE step of
    # dt 2 / 800 substeps; here the same 800 substeps of 0.0025 s run as 40 steps of 20 (same arithmetic)
    n = 'clipped_plane_%s' % mode
    if want(n):
      c = text_format.Parse(pt.CapsuleClippedPlaneTest._CONFIG, brax.Config())
      assert abs(c.dt / c.substeps - 0.0025) < 1e-9
      c.dt, c.substeps = 0.05, 20
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 40)
    # a tilted clipped plane: a capsule (two ends) lands on it, rolls down and leaves over the lower edge, a ball
    # rolls down it and leaves over the side; both end on the ground (the side-plane clipping opens the contacts
    # while the bodies are still above the plane: an end that re-enters the rectangle from below would be
    # ejected by the reference's one-sided penetration -- :776 -- a launch this scene avoids)
    n = 'clipped_ramp_%s' % mode
    if want(n):
      c = text_format.Parse('''
        dt: 0.02 substeps: 8 friction: 0.5 elasticity: 0.2
        gravity { z: -9.8 }
        bodies {
          name: "cap" mass: 1
          colliders { rotation { x: 90 } capsule { radius: .15 length: 1.0 } }
          inertia { x: 0.1 y: 0.1 z: 0.1 }
        }
        bodies {
          name: "ball" mass: 0.5
          colliders { sphere { radius: .2 } }
          inertia { x: 0.01 y: 0.01 z: 0.01 }
        }
        bodies {
          name: "ramp" mass: 1
          colliders { rotation { y: 15 } position { z: 1.0 } clipped_plane { halfsize_x: 1.5 halfsize_y: 0.8 } }
          frozen { all: true }
        }
        bodies { name: "Ground" frozen: { all: true } colliders { plane {}}}
        defaults {
          qps { name: "cap" pos { x: 0.7 y: 0.1 z: 1.2 } vel { z: -0.5 } ang { z: 0.2 } }
          qps { name: "ball" pos { x: -0.8 y: -0.3 z: 1.9 } vel { y: -0.6 } }
          qps { name: "ramp" pos { x: 0 } }
        }
      ''', brax.Config())
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 90)
    # HeightMapTest.test_box_stays_on_heightmap (physics_test.py:229-258): a box falls onto the bottom left of a
    # 3 x 3 height map -- the config AS IS except dt: the reference takes ONE step of dt 2 / 1000 substeps; here
    # the same 1000 substeps of 0.002 s run as 50 steps of 20
    n = 'heightmap_%s' % mode
    if want(n):
      c = text_format.Parse(pt.HeightMapTest._CONFIG, brax.Config())
      assert abs(c.dt / c.substeps - 0.002) < 1e-9
      c.dt, c.substeps = 0.04, 20
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 50)
    # a 5 x 5 terrain on a tilted frozen body: a spinning box lands on a slope, slides / tumbles across cells
    # (both triangles of a square, several squares), a second box rests in a hollow
    n = 'terrain_%s' % mode
    if want(n):
      c = text_format.Parse('''
        dt: 0.02 substeps: 8 friction: 0.7 elasticity: 0.1
        gravity { z: -9.8 }
        bodies {
          name: "crate" mass: 1
          colliders { box { halfsize { x: 0.3 y: 0.2 z: 0.25 } } }
          inertia { x: 0.1 y: 0.1 z: 0.1 }
        }
        bodies {
          name: "brick" mass: 2
          colliders { rotation { z: 30 } box { halfsize { x: 0.25 y: 0.25 z: 0.15 } } }
          inertia { x: 0.2 y: 0.2 z: 0.2 }
        }
        bodies {
          name: "terrain"
          frozen: { all: true }
          colliders {
            heightMap {
              size: 8
              data: [0.0, 0.3, 0.8, 0.4, 0.1,
                     0.2, 0.1, 0.5, 0.9, 0.3,
                     0.6, 0.0, 0.2, 0.5, 0.7,
                     0.9, 0.4, 0.0, 0.1, 0.5,
                     1.2, 0.8, 0.3, 0.0, 0.2]
            }
          }
        }
        collide_include { first: "crate" second: "terrain" }
        collide_include { first: "brick" second: "terrain" }
        defaults {
          qps { name: "crate" pos { x: 2.6 y: -2.2 z: 1.8 } vel { x: 0.8 y: -0.5 z: -0.5 } ang { x: 0.6 y: -0.4 z: 0.3 } }
          qps { name: "brick" pos { x: 5.0 y: -5.2 z: 1.0 } }
          qps { name: "terrain" pos { x: 0.0 y: 0.0 z: 0.0 } rot { x: 4 y: -3 } }
        }
      ''', brax.Config())
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 90)
    n = 'ball_plane_%s_e50' % mode
    if want(n):
      c = text_format.Parse('''
        dt: 0.01 substeps: 10 friction: 0.5 elasticity: 0.5
        gravity { z: -9.8 }
        bodies {
          name: "ball" mass: 1
          colliders { capsule { radius: .5 length: 1.0 } }
          inertia { x: 1 y: 1 z: 1 }
        }
        bodies { name: "Ground" frozen: { all: true } colliders { plane {}}}
        defaults { qps { name: "ball" pos { z: 1.0 } vel { x: 1.0 z: -3.0 } ang { y: 1.0 } } }
      ''', brax.Config())
      if mode == 'spring':
        c.dynamics_mode = 'legacy_spring'
      roll(n, c, 0, 120)


if __name__ == '__main__':
  main(sys.argv[1:])
