/*
 * brax_step_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * A plain-C, single-environment, straight-line restatement of the reference's
 * Brax v1 `System.step` (Prasaya/brax, /root/reference), compiled twice:
 *   -DBXO_DOUBLE : float64, pinned against the reference itself run on its
 *                  numpy backend (tests/golden/, made by oracle/gen_golden.py)
 *                  to <= 1e-12, and against the reference's own known-answer
 *                  tests (geometry_test.py:214-272, colliders_test.py:54-135,
 *                  physics_test.py behavioural answers);
 *   (default)    : float32, the arithmetic the reference runs in under JAX
 *                  defaults -- the checker for the CUDA path and the timed CPU
 *                  baseline ("port") in bench.py.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 *
 * Each function cites the reference file:line it follows.  The reference is
 * Apache-2.0 Python; nothing here is copied text -- it is the same arithmetic
 * written as scalar C in the reference's operation order.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
static int omp_get_max_threads_wrap(void) { return omp_get_max_threads(); }
#endif

#ifdef BXO_COUNT
/* operation-counting build (C++): see oracle/flopcount.hpp */
#include "flopcount.hpp"
double bxo_counters[CNT_N];
typedef CR R;
#define FN(n) n##_cnt
#define SQRT cr_sqrt
#define SIN cr_sin
#define COS cr_cos
#define ATAN2 cr_atan2
#define ACOS cr_acos
#define FABS cr_fabs
#define _Thread_local thread_local
#elif defined(BXO_DOUBLE)
typedef double R;
#define FN(n) n##_f64
#define SQRT sqrt
#define SIN sin
#define COS cos
#define ATAN2 atan2
#define ACOS acos
#define FABS fabs
#define FLOOR floor
#else
typedef float R;
#define FN(n) n##_f32
#define SQRT sqrtf
#define SIN sinf
#define COS cosf
#define ATAN2 atan2f
#define ACOS acosf
#define FABS fabsf
#define FLOOR floorf
#endif

#define C(x) ((R)(x))

enum { MODE_PBD = 0, MODE_SPRING = 1 };
enum { J_REV = 0, J_SPH = 1, J_SREV = 2, J_SUNI = 3, J_SSPH = 4 };
enum { A_TORQUE = 0, A_ANGLE = 1 };
enum { F_THRUSTER = 0, F_TWISTER = 1 };
enum { P_CAPSULE_PLANE = 0, P_CAPSULE_CAPSULE = 1, P_BOX_PLANE = 2, P_CAPSULE_BOX = 3, P_CAPSULE_CLIPPED_PLANE = 4,
       P_BOX_HEIGHTMAP = 5 };

#define MAXB 64
#define MAXK 128
#define MAXG 8

typedef struct {
  R mass, inv_inertia[3], pos_mask[3], rot_mask[3];
} obody;

typedef struct {
  int32_t kind, body_p, body_c, group;
  R off_p[3], off_c[3], axis_p[9], axis_c[9], limit[6];
  R angular_damping, scale_pos, scale_ang;
  R stiffness, spring_damping, limit_strength;
} ojoint;

typedef struct {
  int32_t kind, joint, group, act_index[3];
  R strength;
} oactuator;

typedef struct {
  int32_t kind, body, act_index[3];
  R strength;
} oforce;

typedef struct {
  int32_t kind, group, two_way, body_a, body_b, n_contacts;
  R friction, elasticity, radius_a, radius_b;
  R pts[8][3];
  /* capsule_box (capsule_mesh against geometry.TriangulatedBox): the 12 triangles of the box
   * in the body-b frame (BaseMesh.faces after its winding fix, geometry.py:146-156) and their
   * normals (BoxMesh, geometry.py:162-189) */
  R tri[12][3][3], tri_normal[12][3];
  /* box_heightmap: geometry.HeightMap (geometry.py:292-308) -- hm_size x hm_size heights, row-major */
  const R* hm_height;
  int32_t hm_size;
  R hm_cell_size;
} opair;

typedef struct {
  int32_t mode, substeps, nb, nj, na, nf, np, ngroups, act_size;
  int32_t safe_norm_jax; /* 1: jnp zero rule (jumpy.py:184-189); 0: numpy */
  R dt, gravity[3], vel_damp, ang_damp, erp, collide_scale, vth;
  const obody* bodies;
  const ojoint* joints;
  const oactuator* actuators;
  const oforce* forces;
  const opair* pairs;
} osys;

/* state of one env: QP (base.py:75-133) */
typedef struct {
  R pos[MAXB][3], rot[MAXB][4], vel[MAXB][3], ang[MAXB][3];
} oqp;

typedef struct { R vel[MAXB][3], ang[MAXB][3]; } op;  /* P  base.py:50-72 */
typedef struct { R pos[MAXB][3], rot[MAXB][4]; } oq;  /* Q  base.py:28-47 */

/* ---- small vector kernel (brax/math.py) -------------------------------- */
static inline R dot3(const R* a, const R* b) {
  /* numpy/jnp dot of 3-vectors: sequential sum */
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}
static inline void cross3(const R* a, const R* b, R* o) {
  R x = a[1] * b[2] - a[2] * b[1];
  R y = a[2] * b[0] - a[0] * b[2];
  R z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
/* math.py:25-40 rotate: 2(u.v)u + (s^2 - u.u)v + 2s(u x v) */
static inline void rotate(const R* v, const R* q, R* o) {
  R s = q[0];
  const R* u = q + 1;
  R uv = dot3(u, v), uu = dot3(u, u);
  R cr[3];
  cross3(u, v, cr);
  R r0 = C(2) * (uv * u[0]) + (s * s - uu) * v[0];
  R r1 = C(2) * (uv * u[1]) + (s * s - uu) * v[1];
  R r2 = C(2) * (uv * u[2]) + (s * s - uu) * v[2];
  o[0] = r0 + C(2) * s * cr[0];
  o[1] = r1 + C(2) * s * cr[1];
  o[2] = r2 + C(2) * s * cr[2];
}
/* math.py:130-145 */
static inline void quat_mul(const R* u, const R* v, R* o) {
  R w = u[0] * v[0] - u[1] * v[1] - u[2] * v[2] - u[3] * v[3];
  R x = u[0] * v[1] + u[1] * v[0] + u[2] * v[3] - u[3] * v[2];
  R y = u[0] * v[2] - u[1] * v[3] + u[2] * v[0] + u[3] * v[1];
  R z = u[0] * v[3] + u[1] * v[2] - u[2] * v[1] + u[3] * v[0];
  o[0] = w; o[1] = x; o[2] = y; o[3] = z;
}
/* math.py:148-170 */
static inline void vec_quat_mul(const R* u, const R* v, R* o) {
  R w = -u[0] * v[1] - u[1] * v[2] - u[2] * v[3];
  R x = u[0] * v[0] + u[1] * v[3] - u[2] * v[2];
  R y = -u[0] * v[3] + u[1] * v[0] + u[2] * v[1];
  R z = u[0] * v[2] - u[1] * v[1] + u[2] * v[0];
  o[0] = w; o[1] = x; o[2] = y; o[3] = z;
}
/* math.py:116-127 */
static inline R signed_angle(const R* axis, const R* ref_p, const R* ref_c) {
  R cr[3];
  cross3(ref_p, ref_c, cr);
  return ATAN2(dot3(cr, axis), dot3(ref_p, ref_c));
}
/* math.py:173-187 */
static inline void quat_rot_axis(const R* axis, R angle, R* q) {
  R s = SIN(angle / C(2));
  q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
  q[0] = COS(angle / C(2));
}
/* jumpy.py:170-192 safe_norm.  Under jnp: 0 if allclose(x, 0) i.e. every
 * |x_i| <= 1e-8, else ||x||; under numpy: plain norm. */
static inline R safe_norm3(const osys* s, const R* x) {
  if (s->safe_norm_jax && FABS(x[0]) <= C(1e-8) && FABS(x[1]) <= C(1e-8) &&
      FABS(x[2]) <= C(1e-8))
    return C(0);
  return SQRT(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
}
static inline R clipr(R x, R lo, R hi) { /* np.clip = minimum(maximum(x,lo),hi) */
  R y = x < lo ? lo : x;
  return y > hi ? hi : y;
}
static inline R signr(R x) { return x > 0 ? C(1) : (x < 0 ? C(-1) : C(0)); }

/* ---- joint axis/angle --------------------------------------------------- */
/* Revolute.axis_angle joints.py:311-319 (= spring_joints.py:150-157) */
static void axis_angle_revolute(const ojoint* j, const R* rot_p, const R* rot_c,
                                R axis[3][3], R angle[3]) {
  R ref_p[3], ref_c[3];
  rotate(j->axis_p + 0, rot_p, axis[0]);
  rotate(j->axis_p + 6, rot_p, ref_p);
  rotate(j->axis_c + 6, rot_c, ref_c);
  angle[0] = signed_angle(axis[0], ref_p, ref_c);
}

/* Spherical.axis_angle joints.py:388-415 (= spring_joints.py:212-224 for the
 * Universal joint's first two dofs, :276-303 for the spring Spherical). */
static void axis_angle_spherical(const osys* s, const ojoint* j, const R* rot_p,
                                 const R* rot_c, R axis[3][3], R angle[3]) {
  R a1p[3], a2p[3], a1c[3], a2c[3], a3c[3];
  rotate(j->axis_p + 0, rot_p, a1p);
  rotate(j->axis_p + 3, rot_p, a2p);
  rotate(j->axis_c + 0, rot_c, a1c);
  rotate(j->axis_c + 3, rot_c, a2c);
  rotate(j->axis_c + 6, rot_c, a3c);
  R lon[3];
  cross3(a3c, a1p, lon);
  R nl = C(1e-10) + safe_norm3(s, lon);
  for (int k = 0; k < 3; k++) lon[k] = lon[k] / nl;
  R psi = signed_angle(a1p, a2p, lon);
  R d1 = dot3(a1p, a1c), d2 = dot3(a1p, a2c);
  R xz[3];
  for (int k = 0; k < 3; k++) xz[k] = d1 * a1c[k] + d2 * a2c[k];
  R nx = C(1e-10) + safe_norm3(s, xz);
  for (int k = 0; k < 3; k++) xz[k] = xz[k] / nx;
  R cosang = dot3(xz, a1p);
  R theta = ACOS(clipr(cosang, C(-1), C(1))) * signr(dot3(a1p, a3c));
  R ycn[3] = {-a3c[0], -a3c[1], -a3c[2]};
  R phi = signed_angle(ycn, a2c, lon);
  for (int k = 0; k < 3; k++) {
    axis[0][k] = a1p[k]; axis[1][k] = a2c[k]; axis[2][k] = a3c[k];
  }
  angle[0] = psi; angle[1] = theta; angle[2] = phi;
}

static int joint_dof(const ojoint* j) {
  switch (j->kind) {
    case J_REV: case J_SREV: return 1;
    case J_SUNI: return 2;
    default: return 3;
  }
}

static void joint_axis_angle(const osys* s, const ojoint* j, const R* rot_p,
                             const R* rot_c, R axis[3][3], R angle[3]) {
  if (j->kind == J_REV || j->kind == J_SREV)
    axis_angle_revolute(j, rot_p, rot_c, axis, angle);
  else
    axis_angle_spherical(s, j, rot_p, rot_c, axis, angle);
}

static void zero_p(const osys* s, op* p) {
  for (int b = 0; b < s->nb; b++)
    for (int k = 0; k < 3; k++) p->vel[b][k] = p->ang[b][k] = C(0);
}
static void add_p(const osys* s, op* a, const op* b) {
  for (int i = 0; i < s->nb; i++)
    for (int k = 0; k < 3; k++) {
      a->vel[i][k] = a->vel[i][k] + b->vel[i][k];
      a->ang[i][k] = a->ang[i][k] + b->ang[i][k];
    }
}

/* ---- actuators: actuators.py:52-112 ------------------------------------- */
/* Sum over actuator groups of Actuator.apply; within a group the
 * segment_sum runs over concat(parents, children) (actuators.py:68-70). */
static void actuators_apply(const osys* s, const oqp* qp, const R* act, op* out) {
  zero_p(s, out);
  int i0 = 0;
  while (i0 < s->na) {
    int g = s->actuators[i0].group, i1 = i0;
    while (i1 < s->na && s->actuators[i1].group == g) i1++;
    op grp;
    zero_p(s, &grp);
    R dang_p[64][3], dang_c[64][3];
    for (int i = i0; i < i1; i++) {
      const oactuator* a = &s->actuators[i];
      const ojoint* j = &s->joints[a->joint];
      const obody* bp = &s->bodies[j->body_p];
      const obody* bc = &s->bodies[j->body_c];
      int dof = joint_dof(j);
      R axis[3][3], angle[3], tq[3] = {0, 0, 0};
      joint_axis_angle(s, j, qp->rot[j->body_p], qp->rot[j->body_c], axis, angle);
      for (int d = 0; d < dof; d++) {
        /* take(act, act_index, clip) * act_mask  (actuators.py:64) */
        int idx = a->act_index[d];
        R m = idx >= 0 ? C(1) : C(0);
        if (idx < 0) idx = 0;
        if (idx > s->act_size - 1) idx = s->act_size - 1;
        R av = act[idx] * m;
        R t;
        if (a->kind == A_TORQUE) { /* actuators.py:98-112 */
          t = av * a->strength * C(-1);
          if (angle[d] < j->limit[2 * d]) t = C(0);
          if (angle[d] > j->limit[2 * d + 1]) t = C(0);
        } else { /* Angle actuators.py:79-91 */
          R tgt = clipr(av * C(3.141592653589793) / C(180), j->limit[2 * d],
                        j->limit[2 * d + 1]);
          t = (tgt - angle[d]) * a->strength;
        }
        for (int k = 0; k < 3; k++) tq[k] = tq[k] + axis[d][k] * t;
      }
      for (int k = 0; k < 3; k++) {
        if (a->kind == A_TORQUE) {
          dang_p[i - i0][k] = bp->inv_inertia[k] * tq[k];
          dang_c[i - i0][k] = -bc->inv_inertia[k] * tq[k];
        } else {
          dang_p[i - i0][k] = -bp->inv_inertia[k] * tq[k];
          dang_c[i - i0][k] = bc->inv_inertia[k] * tq[k];
        }
      }
    }
    for (int i = i0; i < i1; i++) {
      int b = s->joints[s->actuators[i].joint].body_p;
      for (int k = 0; k < 3; k++) grp.ang[b][k] += dang_p[i - i0][k];
    }
    for (int i = i0; i < i1; i++) {
      int b = s->joints[s->actuators[i].joint].body_c;
      for (int k = 0; k < 3; k++) grp.ang[b][k] += dang_c[i - i0][k];
    }
    add_p(s, out, &grp);
    i0 = i1;
  }
}

/* ---- forces: forces.py:44-107 ------------------------------------------- */
static void forces_apply(const osys* s, const R* act, op* out) {
  zero_p(s, out);
  for (int kind = 0; kind < 2; kind++) { /* thrusters group, then twisters */
    op grp;
    zero_p(s, &grp);
    int any = 0;
    for (int i = 0; i < s->nf; i++) {
      const oforce* f = &s->forces[i];
      if (f->kind != kind) continue;
      any = 1;
      for (int k = 0; k < 3; k++) {
        int idx = f->act_index[k];
        if (idx < 0) idx = 0;
        if (idx > s->act_size - 1) idx = s->act_size - 1;
        R v = act[idx] * f->strength / s->bodies[f->body].mass;
        if (kind == F_THRUSTER) grp.vel[f->body][k] += v;
        else grp.ang[f->body][k] += v;
      }
    }
    if (any) add_p(s, out, &grp);
  }
}

/* ---- Joint.damp joints.py:103-128 (pbd joints only) ---------------------- */
static void joints_damp(const osys* s, const oqp* qp, op* out) {
  zero_p(s, out);
  int i0 = 0;
  while (i0 < s->nj) {
    int g = s->joints[i0].group, i1 = i0;
    while (i1 < s->nj && s->joints[i1].group == g) i1++;
    op grp;
    zero_p(s, &grp);
    if (s->joints[i0].kind == J_REV || s->joints[i0].kind == J_SPH) {
      R tq[64][3];
      for (int i = i0; i < i1; i++) {
        const ojoint* j = &s->joints[i];
        for (int k = 0; k < 3; k++)
          tq[i - i0][k] = C(-1) * j->angular_damping *
                          (qp->ang[j->body_p][k] - qp->ang[j->body_c][k]);
      }
      for (int i = i0; i < i1; i++) {
        const ojoint* j = &s->joints[i];
        for (int k = 0; k < 3; k++)
          grp.ang[j->body_p][k] += s->bodies[j->body_p].inv_inertia[k] * tq[i - i0][k];
      }
      for (int i = i0; i < i1; i++) {
        const ojoint* j = &s->joints[i];
        for (int k = 0; k < 3; k++)
          grp.ang[j->body_c][k] += -s->bodies[j->body_c].inv_inertia[k] * tq[i - i0][k];
      }
    }
    add_p(s, out, &grp);
    i0 = i1;
  }
}

/* ---- integrator: integrators.py -------------------------------------- */
/* Euler.update acc branch, integrators.py:87-95 */
static void update_acc(const osys* s, oqp* qp, const op* dp) {
  for (int b = 0; b < s->nb; b++) {
    const obody* bd = &s->bodies[b];
    for (int k = 0; k < 3; k++) {
      R v = s->vel_damp * qp->vel[b][k];
      v = v + (dp->vel[b][k] + s->gravity[k]) * s->dt;
      qp->vel[b][k] = v * bd->pos_mask[k];
      R a = s->ang_damp * qp->ang[b][k];
      a = a + dp->ang[b][k] * s->dt;
      qp->ang[b][k] = a * bd->rot_mask[k];
    }
  }
}
/* Euler.update vel branch, integrators.py:97-101 */
static void update_vel(const osys* s, oqp* qp, const op* dp) {
  for (int b = 0; b < s->nb; b++)
    for (int k = 0; k < 3; k++) {
      qp->vel[b][k] = (qp->vel[b][k] + dp->vel[b][k]) * s->bodies[b].pos_mask[k];
      qp->ang[b][k] = (qp->ang[b][k] + dp->ang[b][k]) * s->bodies[b].rot_mask[k];
    }
}
/* Euler.update pos branch, integrators.py:103-110 (quat_mask = [1, rot_mask]) */
static void update_pos(const osys* s, oqp* qp, const oq* dq) {
  for (int b = 0; b < s->nb; b++) {
    const obody* bd = &s->bodies[b];
    for (int k = 0; k < 3; k++)
      qp->pos[b][k] = qp->pos[b][k] + dq->pos[b][k] * bd->pos_mask[k];
    qp->rot[b][0] = qp->rot[b][0] + dq->rot[b][0] * C(1);
    for (int k = 0; k < 3; k++)
      qp->rot[b][k + 1] = qp->rot[b][k + 1] + dq->rot[b][k + 1] * bd->rot_mask[k];
  }
}
/* Euler.kinetic, integrators.py:50-68 */
static void kinetic(const osys* s, oqp* qp) {
  for (int b = 0; b < s->nb; b++) {
    const obody* bd = &s->bodies[b];
    R aq[4], qm[4];
    for (int k = 0; k < 3; k++)
      qp->pos[b][k] = qp->pos[b][k] + qp->vel[b][k] * s->dt * bd->pos_mask[k];
    aq[0] = C(0) * C(0.5) * s->dt;
    for (int k = 0; k < 3; k++)
      aq[k + 1] = (qp->ang[b][k] * bd->rot_mask[k]) * C(0.5) * s->dt;
    quat_mul(aq, qp->rot[b], qm);
    R r[4];
    for (int k = 0; k < 4; k++) r[k] = qp->rot[b][k] + qm[k];
    R n = SQRT(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
    for (int k = 0; k < 4; k++) qp->rot[b][k] = r[k] / n;
  }
}
/* Euler.velocity_projection, integrators.py:122-146 */
static void velocity_projection(const osys* s, oqp* qp, const oqp* prev) {
  for (int b = 0; b < s->nb; b++) {
    const obody* bd = &s->bodies[b];
    R* r = qp->rot[b];
    R n = SQRT(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
    R nr[4];
    for (int k = 0; k < 4; k++) nr[k] = r[k] / n;
    for (int k = 0; k < 3; k++)
      qp->vel[b][k] = ((qp->pos[b][k] - prev->pos[b][k]) / s->dt) * bd->pos_mask[k];
    /* relative_quat(q1=prev, q2=new) = quat_mul(new, quat_inv(prev)) */
    R inv[4] = {prev->rot[b][0] * C(1), prev->rot[b][1] * C(-1),
                prev->rot[b][2] * C(-1), prev->rot[b][3] * C(-1)};
    R dq[4];
    quat_mul(nr, inv, dq);
    R sg = dq[0] >= C(0) ? C(1) : C(-1);
    for (int k = 0; k < 3; k++) {
      R a = C(2) * dq[k + 1] / s->dt;
      R scale = sg * bd->rot_mask[k];
      qp->ang[b][k] = scale * a * bd->rot_mask[k];
    }
    for (int k = 0; k < 4; k++) r[k] = nr[k];
  }
}

/* ---- pbd joints: joints.py ------------------------------------------------ */
typedef struct { R pos[3], rot[4]; } dq1;

/* Joint.apply_angle_update joints.py:130-152; adds into dp, dc */
static void angle_update(const osys* s, const ojoint* j, const R* rot_p,
                         const R* rot_c, const R* dq, dq1* dp, dq1* dc) {
  const obody* bp = &s->bodies[j->body_p];
  const obody* bc = &s->bodies[j->body_c];
  R th = safe_norm3(s, dq);
  R n[3], in_p[3], in_c[3], p[3];
  for (int k = 0; k < 3; k++) n[k] = dq[k] / (th + C(1e-6));
  for (int k = 0; k < 3; k++) { in_p[k] = bp->inv_inertia[k] * n[k]; in_c[k] = bc->inv_inertia[k] * n[k]; }
  R w1 = dot3(n, in_p), w2 = dot3(n, in_c);
  R dl = -th / (w1 + w2 + C(1e-6));
  for (int k = 0; k < 3; k++) p[k] = -dl * n[k];
  R ip[3], ic[3], qp_[4], qc_[4];
  for (int k = 0; k < 3; k++) { ip[k] = bp->inv_inertia[k] * p[k]; ic[k] = bc->inv_inertia[k] * p[k]; }
  vec_quat_mul(ip, rot_p, qp_);
  vec_quat_mul(ic, rot_c, qc_);
  for (int k = 0; k < 4; k++) {
    dp->rot[k] = j->scale_ang * (C(.5) * qp_[k]);
    dc->rot[k] = j->scale_ang * (C(-.5) * qc_[k]);
  }
  for (int k = 0; k < 3; k++) dp->pos[k] = dc->pos[k] = C(0);
}

/* Joint.apply_position_update joints.py:154-195 */
static void position_update(const osys* s, const ojoint* j, const oqp* qp,
                            const R* wp, const R* wc, dq1* dp, dq1* dc) {
  const obody* bp = &s->bodies[j->body_p];
  const obody* bc = &s->bodies[j->body_c];
  const R* rot_p = qp->rot[j->body_p];
  const R* rot_c = qp->rot[j->body_c];
  R dx[3], rp[3], rc[3], n[3];
  for (int k = 0; k < 3; k++) {
    dx[k] = wp[k] - wc[k];
    rp[k] = wp[k] - qp->pos[j->body_p][k];
    rc[k] = wc[k] - qp->pos[j->body_c][k];
  }
  R c = safe_norm3(s, dx);
  for (int k = 0; k < 3; k++) n[k] = dx[k] / (c + C(1e-6));
  R cr1[3], cr2[3], t1[3], t2[3];
  cross3(rp, n, cr1);
  for (int k = 0; k < 3; k++) t1[k] = bp->inv_inertia[k] * cr1[k];
  R w1 = (C(1) / bp->mass) + dot3(cr1, t1);
  cross3(rc, n, cr2);
  for (int k = 0; k < 3; k++) t2[k] = bc->inv_inertia[k] * cr2[k];
  R w2 = (C(1) / bc->mass) + dot3(cr2, t2);
  R dl = -c / (w1 + w2 + C(1e-6));
  R p[3], cp[3], cc[3], ip[3], ic[3], qp_[4], qc_[4];
  for (int k = 0; k < 3; k++) p[k] = dl * n[k];
  cross3(rp, p, cp);
  cross3(rc, p, cc);
  for (int k = 0; k < 3; k++) { ip[k] = bp->inv_inertia[k] * cp[k]; ic[k] = bc->inv_inertia[k] * cc[k]; }
  vec_quat_mul(ip, rot_p, qp_);
  vec_quat_mul(ic, rot_c, qc_);
  for (int k = 0; k < 3; k++) {
    dp->pos[k] = j->scale_pos * (p[k] / bp->mass);
    dc->pos[k] = j->scale_pos * (-p[k] / bc->mass);
  }
  for (int k = 0; k < 4; k++) {
    dp->rot[k] = j->scale_pos * (C(.5) * qp_[k]);
    dc->rot[k] = j->scale_pos * (C(-.5) * qc_[k]);
  }
}

static void to_world_pos(const oqp* qp, int b, const R* off, R* out) {
  R r[3];
  rotate(off, qp->rot[b], r);
  for (int k = 0; k < 3; k++) out[k] = qp->pos[b][k] + r[k];
}

/* Revolute.apply_reduced joints.py:270-309 */
static void revolute_apply(const osys* s, const ojoint* j, const oqp* qp,
                           dq1* dp, dq1* dc) {
  const R* rot_p = qp->rot[j->body_p];
  const R* rot_c = qp->rot[j->body_c];
  R wp[3], wc[3];
  to_world_pos(qp, j->body_p, j->off_p, wp);
  to_world_pos(qp, j->body_c, j->off_c, wc);
  position_update(s, j, qp, wp, wc, dp, dc);
  R axis[3], ref_p[3], ref_c[3], axis_c[3], dq_1[3], dq_2[3], fix[4], n1[3];
  rotate(j->axis_p + 0, rot_p, axis);
  rotate(j->axis_p + 6, rot_p, ref_p);
  rotate(j->axis_c + 6, rot_c, ref_c);
  R psi = signed_angle(axis, ref_p, ref_c);
  rotate(j->axis_c + 0, rot_c, axis_c);
  cross3(axis, axis_c, dq_1);
  R ph = clipr(psi, j->limit[0], j->limit[1]);
  quat_rot_axis(axis, ph, fix);
  rotate(ref_p, fix, n1);
  cross3(n1, ref_c, dq_2);
  dq1 p1, c1, p2, c2;
  angle_update(s, j, rot_p, rot_c, dq_1, &p1, &c1);
  angle_update(s, j, rot_p, rot_c, dq_2, &p2, &c2);
  for (int k = 0; k < 3; k++) {
    dp->pos[k] = dp->pos[k] + (p1.pos[k] + p2.pos[k]);
    dc->pos[k] = dc->pos[k] + (c1.pos[k] + c2.pos[k]);
  }
  for (int k = 0; k < 4; k++) {
    dp->rot[k] = dp->rot[k] + (p1.rot[k] + p2.rot[k]);
    dc->rot[k] = dc->rot[k] + (c1.rot[k] + c2.rot[k]);
  }
}

/* limit_angle closure of Spherical.apply_reduced, joints.py:343-354 */
static void limit_angle(const osys* s, const ojoint* j, const R* rot_p,
                        const R* rot_c, const R* n, const R* n_1, const R* n_2,
                        int lim, dq1* dp, dq1* dc) {
  R ph = signed_angle(n, n_1, n_2);
  R mask = ph < j->limit[2 * lim] ? C(1) : C(0);
  if (ph > j->limit[2 * lim + 1]) mask = C(1);
  ph = clipr(ph, j->limit[2 * lim], j->limit[2 * lim + 1]);
  R fix[4], n1[3], dq[3];
  quat_rot_axis(n, ph, fix);
  rotate(n_1, fix, n1);
  cross3(n1, n_2, dq);
  for (int k = 0; k < 3; k++) dq[k] = dq[k] * mask;
  angle_update(s, j, rot_p, rot_c, dq, dp, dc);
}

/* Spherical.apply_reduced joints.py:332-386 */
static void spherical_apply(const osys* s, const ojoint* j, const oqp* qp,
                            dq1* dp, dq1* dc) {
  const R* rot_p = qp->rot[j->body_p];
  const R* rot_c = qp->rot[j->body_c];
  R wp[3], wc[3];
  to_world_pos(qp, j->body_p, j->off_p, wp);
  to_world_pos(qp, j->body_c, j->off_c, wc);
  position_update(s, j, qp, wp, wc, dp, dc);
  R a1p[3], a2p[3], a1c[3], a2c[3], a3c[3];
  rotate(j->axis_p + 0, rot_p, a1p);
  rotate(j->axis_p + 3, rot_p, a2p);
  rotate(j->axis_c + 0, rot_c, a1c);
  rotate(j->axis_c + 3, rot_c, a2c);
  rotate(j->axis_c + 6, rot_c, a3c);
  R lon[3], xz[3], a2n[3];
  cross3(a3c, a1p, lon);
  R nl = C(1e-6) + safe_norm3(s, lon);
  for (int k = 0; k < 3; k++) lon[k] = lon[k] / nl;
  R d1 = dot3(a1p, a1c), d2 = dot3(a1p, a2c);
  for (int k = 0; k < 3; k++) xz[k] = d1 * a1c[k] + d2 * a2c[k];
  R nx = C(1e-6) + safe_norm3(s, xz);
  for (int k = 0; k < 3; k++) xz[k] = xz[k] / nx;
  cross3(xz, a1p, a2n);
  R n2 = C(1e-6) + safe_norm3(s, a2n);
  for (int k = 0; k < 3; k++) a2n[k] = a2n[k] / n2;
  R sg = signr(dot3(a1p, a3c));
  R nb[3], yc[3];
  for (int k = 0; k < 3; k++) { nb[k] = -a2n[k] * sg; yc[k] = -(-a3c[k]); }
  dq1 p1, c1, p2, c2, p3, c3;
  limit_angle(s, j, rot_p, rot_c, a1p, a2p, lon, 0, &p1, &c1);
  limit_angle(s, j, rot_p, rot_c, nb, a1p, xz, 1, &p2, &c2);
  limit_angle(s, j, rot_p, rot_c, yc, lon, a2c, 2, &p3, &c3);
  /* dq_p += dq_p_ang_1 + dq_p_ang_2 + dq_p_ang_3  -> dq_p + ((a1 + a2) + a3) */
  for (int k = 0; k < 3; k++) {
    dp->pos[k] = dp->pos[k] + ((p1.pos[k] + p2.pos[k]) + p3.pos[k]);
    dc->pos[k] = dc->pos[k] + ((c1.pos[k] + c2.pos[k]) + c3.pos[k]);
  }
  for (int k = 0; k < 4; k++) {
    dp->rot[k] = dp->rot[k] + ((p1.rot[k] + p2.rot[k]) + p3.rot[k]);
    dc->rot[k] = dc->rot[k] + ((c1.rot[k] + c2.rot[k]) + c3.rot[k]);
  }
}

static void zero_q(const osys* s, oq* q) {
  for (int b = 0; b < s->nb; b++) {
    for (int k = 0; k < 3; k++) q->pos[b][k] = C(0);
    for (int k = 0; k < 4; k++) q->rot[b][k] = C(0);
  }
}
static void add_q(const osys* s, oq* a, const oq* b) {
  for (int i = 0; i < s->nb; i++) {
    for (int k = 0; k < 3; k++) a->pos[i][k] = a->pos[i][k] + b->pos[i][k];
    for (int k = 0; k < 4; k++) a->rot[i][k] = a->rot[i][k] + b->rot[i][k];
  }
}

/* sum over pbd joint groups of Joint.apply, joints.py:79-100 */
static void joints_apply_pbd(const osys* s, const oqp* qp, oq* out) {
  zero_q(s, out);
  int i0 = 0;
  while (i0 < s->nj) {
    int g = s->joints[i0].group, i1 = i0;
    while (i1 < s->nj && s->joints[i1].group == g) i1++;
    if (s->joints[i0].kind == J_REV || s->joints[i0].kind == J_SPH) {
      oq grp;
      zero_q(s, &grp);
      dq1 dps[64], dcs[64];
      for (int i = i0; i < i1; i++) {
        if (s->joints[i].kind == J_REV)
          revolute_apply(s, &s->joints[i], qp, &dps[i - i0], &dcs[i - i0]);
        else
          spherical_apply(s, &s->joints[i], qp, &dps[i - i0], &dcs[i - i0]);
      }
      for (int i = i0; i < i1; i++) {
        int b = s->joints[i].body_p;
        for (int k = 0; k < 3; k++) grp.pos[b][k] += dps[i - i0].pos[k];
        for (int k = 0; k < 4; k++) grp.rot[b][k] += dps[i - i0].rot[k];
      }
      for (int i = i0; i < i1; i++) {
        int b = s->joints[i].body_c;
        for (int k = 0; k < 3; k++) grp.pos[b][k] += dcs[i - i0].pos[k];
        for (int k = 0; k < 4; k++) grp.rot[b][k] += dcs[i - i0].rot[k];
      }
      add_q(s, out, &grp);
    }
    i0 = i1;
  }
}

/* ---- contact functions: colliders.py / geometry.py ------------------------ */
typedef struct { R pos[3], vel[3], normal[3], pen; } ocontact;

/* geometry.py:360-374 */
static R closest_segment_point_and_dist(const R* a, const R* b, const R* pt, R* p) {
  R ab[3], pa[3], v[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; pa[k] = pt[k] - a[k]; }
  R t = dot3(pa, ab) / (dot3(ab, ab) + C(1e-6));
  t = clipr(t, C(0), C(1));
  for (int k = 0; k < 3; k++) { p[k] = a[k] + t * ab[k]; v[k] = pt[k] - p[k]; }
  return dot3(v, v);
}

/* geometry.py:394-459 closest_segment_to_segment_points */
static void closest_seg_seg(const osys* s, const R* a0, const R* a1, const R* b0,
                            const R* b1, R* best_a, R* best_b) {
  R dir_a[3], dir_b[3], a_mid[3], b_mid[3], trans[3];
  for (int k = 0; k < 3; k++) dir_a[k] = a1[k] - a0[k];
  R len_a = safe_norm3(s, dir_a);
  len_a += C(1e-6) * (len_a == C(0) ? C(1) : C(0));
  for (int k = 0; k < 3; k++) dir_a[k] = dir_a[k] / len_a;
  R half_a = len_a * C(0.5);
  for (int k = 0; k < 3; k++) dir_b[k] = b1[k] - b0[k];
  R len_b = safe_norm3(s, dir_b);
  len_b += C(1e-6) * (len_b == C(0) ? C(1) : C(0));
  for (int k = 0; k < 3; k++) dir_b[k] = dir_b[k] / len_b;
  R half_b = len_b * C(0.5);
  for (int k = 0; k < 3; k++) {
    a_mid[k] = a0[k] + dir_a[k] * half_a;
    b_mid[k] = b0[k] + dir_b[k] * half_b;
    trans[k] = a_mid[k] - b_mid[k];
  }
  R dab = dot3(dir_a, dir_b), dat = dot3(dir_a, trans), dbt = dot3(dir_b, trans);
  R denom = C(1) - dab * dab;
  R ta0 = (-dat + dab * dbt) / (denom + C(1e-6));
  R tb0 = dbt + ta0 * dab;
  R ta = clipr(ta0, -half_a, half_a);
  R tb = clipr(tb0, -half_b, half_b);
  R ba[3], bb[3], na[3], nb[3];
  for (int k = 0; k < 3; k++) {
    ba[k] = a_mid[k] + dir_a[k] * ta;
    bb[k] = b_mid[k] + dir_b[k] * tb;
  }
  R d1 = closest_segment_point_and_dist(a0, a1, bb, na);
  R d2 = closest_segment_point_and_dist(b0, b1, ba, nb);
  for (int k = 0; k < 3; k++) {
    best_a[k] = d1 < d2 ? na[k] : ba[k];
    best_b[k] = d1 < d2 ? bb[k] : nb[k];
  }
}

/* geometry.py:377-391 */
static void closest_segment_point_plane(const R* a, const R* b, const R* p0, const R* n, R* out) {
  R ab[3];
  for (int k = 0; k < 3; k++) ab[k] = b[k] - a[k];
  R d = dot3(p0, n);
  R denom = dot3(n, ab);
  R t = (d - dot3(n, a)) / (denom + C(1e-6));
  t = clipr(t, C(0), C(1));
  for (int k = 0; k < 3; k++) out[k] = a[k] + t * ab[k];
}

/* geometry.py:462-500 closest_triangle_point */
static void closest_triangle_point(const R* p0, const R* p1, const R* p2, const R* pt, R* out) {
  R e0[3], e1[3], d[3], cp[3], df[3];
  for (int k = 0; k < 3; k++) { e0[k] = p1[k] - p0[k]; e1[k] = p2[k] - p0[k]; d[k] = pt[k] - p0[k]; }
  R a = dot3(e0, e0), b = dot3(e0, e1), c = dot3(e1, e1);
  R det = a * c - b * b;
  R u = (c * dot3(e0, d) - b * dot3(e1, d)) / det;
  R v = (-b * dot3(e0, d) + a * dot3(e1, d)) / det;
  int inside = (C(0) <= u) && (u <= C(1)) && (C(0) <= v) && (v <= C(1)) && (u + v <= C(1));
  for (int k = 0; k < 3; k++) { cp[k] = p0[k] + u * e0[k] + v * e1[k]; df[k] = cp[k] - pt[k]; }
  R d0 = dot3(df, df);
  R c1[3], c2[3], c3[3];
  R d1 = closest_segment_point_and_dist(p0, p1, pt, c1);
  int take0 = (d0 < d1) && inside;
  R min_d = take0 ? d0 : d1;
  for (int k = 0; k < 3; k++) out[k] = take0 ? cp[k] : c1[k];
  R d2 = closest_segment_point_and_dist(p1, p2, pt, c2);
  if (d2 < min_d) for (int k = 0; k < 3; k++) out[k] = c2[k];
  min_d = d2 < min_d ? d2 : min_d; /* jp.minimum */
  R d3 = closest_segment_point_and_dist(p2, p0, pt, c3);
  if (d3 < min_d) for (int k = 0; k < 3; k++) out[k] = c3[k];
}

/* geometry.py:503-540 closest_segment_triangle_points */
static void closest_segment_triangle_points(const osys* s, const R* a, const R* b, const R* p0, const R* p1,
                                            const R* p2, const R* tn, R* seg_pt, R* tri_pt) {
  R sp[4][3], tp[4][3], dd[4];
  closest_seg_seg(s, a, b, p0, p1, sp[0], tp[0]);
  closest_seg_seg(s, a, b, p1, p2, sp[1], tp[1]);
  closest_seg_seg(s, a, b, p0, p2, sp[2], tp[2]);
  closest_segment_point_plane(a, b, p0, tn, sp[3]);
  closest_triangle_point(p0, p1, p2, sp[3], tp[3]);
  for (int i = 0; i < 4; i++) {
    R v[3];
    for (int k = 0; k < 3; k++) v[k] = sp[i][k] - tp[i][k];
    dd[i] = dot3(v, v);
  }
  R mn = dd[0];
  for (int i = 1; i < 4; i++) mn = dd[i] < mn ? dd[i] : mn;
  /* mask = (distance == min); the candidates that tie are AVERAGED (sum * mask / sum(mask)) */
  R cnt = C(0);
  for (int k = 0; k < 3; k++) { seg_pt[k] = C(0); tri_pt[k] = C(0); }
  for (int i = 0; i < 4; i++) {
    R m = dd[i] == mn ? C(1) : C(0);
    cnt += m;
    for (int k = 0; k < 3; k++) { seg_pt[k] += sp[i][k] * m; tri_pt[k] += tp[i][k] * m; }
  }
  for (int k = 0; k < 3; k++) { seg_pt[k] = seg_pt[k] / cnt; tri_pt[k] = tri_pt[k] / cnt; }
}

static void world_velocity(const oqp* qp, int b, const R* pos, R* out) {
  R r[3], cr[3];
  for (int k = 0; k < 3; k++) r[k] = pos[k] - qp->pos[b][k];
  cross3(qp->ang[b], r, cr);
  for (int k = 0; k < 3; k++) out[k] = qp->vel[b][k] + cr[k];
}

/* contact_fn dispatch for one pair -> n_contacts rows */
static void pair_contacts(const osys* s, const opair* p, const oqp* qp, ocontact* out) {
  int a = p->body_a, b = p->body_b;
  const R zaxis[3] = {0, 0, 1};
  if (p->kind == P_CAPSULE_PLANE) { /* colliders.py:744-759 */
    for (int c = 0; c < p->n_contacts; c++) {
      R e[3], n[3], r[3], cr[3], d[3];
      rotate(p->pts[c], qp->rot[a], e);
      for (int k = 0; k < 3; k++) e[k] = qp->pos[a][k] + e[k];
      rotate(zaxis, qp->rot[b], n);
      for (int k = 0; k < 3; k++) {
        out[c].pos[k] = e[k] - n[k] * p->radius_a;
        r[k] = out[c].pos[k] - qp->pos[a][k];
      }
      cross3(qp->ang[a], r, cr);
      for (int k = 0; k < 3; k++) {
        out[c].vel[k] = qp->vel[a][k] + cr[k];
        out[c].normal[k] = n[k];
        d[k] = qp->pos[b][k] - out[c].pos[k];
      }
      out[c].pen = dot3(d, n);
    }
  } else if (p->kind == P_BOX_PLANE) { /* colliders.py:667-680, base.py:112-125 */
    for (int c = 0; c < p->n_contacts; c++) {
      R ro[3], cr[3], n[3], d[3];
      rotate(p->pts[c], qp->rot[a], ro);
      cross3(qp->ang[a], ro, cr);
      rotate(zaxis, qp->rot[b], n);
      for (int k = 0; k < 3; k++) {
        out[c].pos[k] = qp->pos[a][k] + ro[k];
        out[c].vel[k] = qp->vel[a][k] + cr[k];
        out[c].normal[k] = n[k];
        d[k] = qp->pos[b][k] - out[c].pos[k];
      }
      out[c].pen = dot3(d, n);
    }
#ifndef BXO_COUNT   /* (no env config has a height map: the operation-counting build leaves it out) */
  } else if (p->kind == P_BOX_HEIGHTMAP) { /* colliders.py:699-742 */
    const int n = p->hm_size;
    const R cell = p->hm_cell_size;
    R qinv[4] = {qp->rot[b][0], -qp->rot[b][1], -qp->rot[b][2], -qp->rot[b][3]}; /* math.inv_rotate */
    for (int c = 0; c < p->n_contacts; c++) {
      R ro[3], cr[3], d[3], lp[3];
      rotate(p->pts[c], qp->rot[a], ro);
      cross3(qp->ang[a], ro, cr);
      for (int k = 0; k < 3; k++) {
        out[c].pos[k] = qp->pos[a][k] + ro[k];
        out[c].vel[k] = qp->vel[a][k] + cr[k];
        d[k] = out[c].pos[k] - qp->pos[b][k];
      }
      rotate(d, qinv, lp);
      R u = lp[0] / cell, v = lp[1] / cell;
      R fu = FLOOR(u), fv = FLOOR(v);
      int iu = (int)fu, iv = (int)fv;
      int lower = (u - fu) + (v - fv) < C(1.);
      R mu = lower ? C(-1.) : C(1.);
      int tu[3], tv[3];
      tu[0] = iu + (lower ? 0 : 1); tu[1] = iu + (lower ? 1 : 0); tu[2] = iu + (lower ? 0 : 1);
      tv[0] = iv + (lower ? 0 : 1); tv[1] = iv + (lower ? 0 : 1); tv[2] = iv + (lower ? 1 : 0);
      R hh[3];
      for (int i = 0; i < 3; i++) { /* height[u, -v]: a negative index wraps once, the rest is clamped (jax gather) */
        int ia = tu[i], ib = -tv[i];
        if (ia < 0) ia += n;
        if (ib < 0) ib += n;
        ia = ia < 0 ? 0 : (ia > n - 1 ? n - 1 : ia);
        ib = ib < 0 ? 0 : (ib > n - 1 ? n - 1 : ib);
        hh[i] = p->hm_height[ia * n + ib];
      }
      R raw[3] = {mu * (hh[1] - hh[0]), mu * (hh[2] - hh[0]), cell}, nl[3], p0[3], dd[3];
      R nrm = safe_norm3(s, raw);
      for (int k = 0; k < 3; k++) nl[k] = raw[k] / nrm;
      p0[0] = (R)tu[0] * cell; p0[1] = (R)tv[0] * cell; p0[2] = hh[0];
      for (int k = 0; k < 3; k++) dd[k] = p0[k] - lp[k];
      out[c].pen = dot3(dd, nl);
      rotate(nl, qp->rot[b], out[c].normal);
    }
#endif
  } else if (p->kind == P_CAPSULE_CLIPPED_PLANE) { /* colliders.py:762-802; plane data in tri[0], tri[1] */
    const R* pl_normal = p->tri[0][0];
    const R* pl_pos = p->tri[0][1];
    const R* pl_x = p->tri[0][2];
    const R* pl_y = p->tri[1][0];
    const R hx = p->tri[1][1][0], hy = p->tri[1][1][1];
    for (int c = 0; c < p->n_contacts; c++) {
      R e[3], n[3], r[3], cr[3], d[3], ppt[3], nx[3], ny[3], nn[3], yn[3], xn[3];
      rotate(p->pts[c], qp->rot[a], e);
      for (int k = 0; k < 3; k++) e[k] = qp->pos[a][k] + e[k];
      rotate(pl_normal, qp->rot[b], n);
      /* orient the normal s.t. it points at the CoM of the capsule (:771-773) */
      R dir = dot3(qp->pos[a], n) > C(0.) ? C(1.) : C(-1.);
      for (int k = 0; k < 3; k++) n[k] = n[k] * dir;
      for (int k = 0; k < 3; k++) {
        out[c].pos[k] = e[k] - n[k] * p->radius_a;
        r[k] = out[c].pos[k] - qp->pos[a][k];
      }
      cross3(qp->ang[a], r, cr);
      rotate(pl_pos, qp->rot[b], ppt);
      for (int k = 0; k < 3; k++) {
        out[c].vel[k] = qp->vel[a][k] + cr[k];
        out[c].normal[k] = n[k];
        ppt[k] = ppt[k] + qp->pos[b][k];
        d[k] = ppt[k] - out[c].pos[k];
      }
      R pen = dot3(d, n);
      /* clip against side planes (:779-798) */
      rotate(pl_x, qp->rot[b], nx);
      rotate(pl_y, qp->rot[b], ny);
      for (int k = 0; k < 3; k++) nn[k] = n[k] * dir;
      cross3(nn, nx, yn);
      cross3(nn, ny, xn);
      for (int k = 0; k < 3; k++) xn[k] = -xn[k];
      int open = 0;
      for (int sd = 0; sd < 4; sd++) {
        const R* dv = sd < 2 ? nx : ny;
        const R* sn = sd < 2 ? xn : yn;
        R hs = sd < 2 ? hx : hy, sg = (sd & 1) ? C(-1.) : C(1.);
        R spt[3], rel[3], snn[3];
        for (int k = 0; k < 3; k++) {
          spt[k] = (sd & 1) ? ppt[k] - dv[k] * hs : ppt[k] + dv[k] * hs;
          rel[k] = out[c].pos[k] - spt[k];
          snn[k] = sg * sn[k];
        }
        if (dot3(rel, snn) > C(1e-6)) open = 1; /* geometry.py:563-565 */
      }
      out[c].pen = open ? C(-1.) : pen;
    }
  } else if (p->kind == P_CAPSULE_BOX) { /* capsule_mesh colliders.py:822-849, one row per triangle */
    R pa[3], ea[3], a0[3], a1[3];
    rotate(p->pts[1], qp->rot[a], pa);
    rotate(p->pts[0], qp->rot[a], ea);
    for (int k = 0; k < 3; k++) {
      pa[k] = qp->pos[a][k] + pa[k];
      a0[k] = pa[k] + ea[k]; a1[k] = pa[k] - ea[k];
    }
    for (int c = 0; c < p->n_contacts; c++) {
      R tn[3], v[3][3], sp[3], tp[3], pv[3], va[3], vb[3];
      rotate(p->tri_normal[c], qp->rot[b], tn);
      for (int i = 0; i < 3; i++) {
        rotate(p->tri[c][i], qp->rot[b], v[i]);
        for (int k = 0; k < 3; k++) v[i][k] = qp->pos[b][k] + v[i][k];
      }
      closest_segment_triangle_points(s, a0, a1, v[0], v[1], v[2], tn, sp, tp);
      for (int k = 0; k < 3; k++) pv[k] = sp[k] - tp[k];
      R dist = safe_norm3(s, pv);
      for (int k = 0; k < 3; k++) {
        out[c].normal[k] = pv[k] / (C(1e-6) + dist);
        out[c].pos[k] = tp[k];
      }
      out[c].pen = p->radius_a - dist;
      world_velocity(qp, a, out[c].pos, va);
      world_velocity(qp, b, out[c].pos, vb);
      for (int k = 0; k < 3; k++) out[c].vel[k] = va[k] - vb[k];
    }
  } else { /* capsule_capsule colliders.py:805-819, _endpoints :661-664 */
    R pa[3], ea[3], pb[3], eb[3], a0[3], a1[3], b0[3], b1[3], ba[3], bb[3];
    rotate(p->pts[1], qp->rot[a], pa);
    rotate(p->pts[0], qp->rot[a], ea);
    rotate(p->pts[3], qp->rot[b], pb);
    rotate(p->pts[2], qp->rot[b], eb);
    for (int k = 0; k < 3; k++) {
      pa[k] = qp->pos[a][k] + pa[k];
      pb[k] = qp->pos[b][k] + pb[k];
      a0[k] = pa[k] + ea[k]; a1[k] = pa[k] - ea[k];
      b0[k] = pb[k] + eb[k]; b1[k] = pb[k] - eb[k];
    }
    closest_seg_seg(s, a0, a1, b0, b1, ba, bb);
    R pv[3], va[3], vb[3];
    for (int k = 0; k < 3; k++) pv[k] = ba[k] - bb[k];
    R dist = safe_norm3(s, pv);
    for (int k = 0; k < 3; k++) {
      out[0].normal[k] = pv[k] / (C(1e-6) + dist);
      out[0].pos[k] = (ba[k] + bb[k]) / C(2);
    }
    out[0].pen = p->radius_a + p->radius_b - dist;
    world_velocity(qp, a, out[0].pos, va);
    world_velocity(qp, b, out[0].pos, vb);
    for (int k = 0; k < 3; k++) out[0].vel[k] = va[k] - vb[k];
  }
}

/* Exported for known-answer tests (geometry_test.py:214-272,
 * colliders_test.py:54-135). */
void FN(bxo_closest_segment_points)(const R* a0, const R* a1, const R* b0,
                                    const R* b1, int jax_norm, R* best_a, R* best_b) {
  osys s;
  memset(&s, 0, sizeof(s));
  s.safe_norm_jax = jax_norm;
  closest_seg_seg(&s, a0, a1, b0, b1, best_a, best_b);
}

/* ---- pbd collision response: colliders.py:306-442 (one-way) and
 *      :495-658 (two-way) ------------------------------------------------- */
static void half_vqm(const obody* bd, const R* r, const R* p, const R* rot, R sgn,
                     R* out4) {
  /* sgn * vec_quat_mul(inertia * cross(r, p), rot) */
  R cr[3], ic[3], q[4];
  cross3(r, p, cr);
  for (int k = 0; k < 3; k++) ic[k] = bd->inv_inertia[k] * cr[k];
  vec_quat_mul(ic, rot, q);
  for (int k = 0; k < 4; k++) out4[k] = sgn * q[k];
}

static R w_of(const obody* bd, const R* r, const R* n) {
  R cr[3], t[3];
  cross3(r, n, cr);
  for (int k = 0; k < 3; k++) t[k] = bd->inv_inertia[k] * cr[k];
  return (C(1) / bd->mass) + dot3(cr, t);
}

/* OneWayCollider._position_contact colliders.py:306-377 */
static R oneway_position(const osys* s, const opair* p, const oqp* qp,
                         const oqp* qprev, const ocontact* ct, dq1* dqa) {
  int a = p->body_a, b = p->body_b;
  const obody* ba = &s->bodies[a];
  R pos_c[3], dx[3], rp[3], n[3];
  for (int k = 0; k < 3; k++) {
    pos_c[k] = ct->pos[k] + ct->normal[k] * ct->pen;
    dx[k] = ct->pos[k] - pos_c[k];
    rp[k] = ct->pos[k] - qp->pos[a][k];
    n[k] = ct->normal[k];
  }
  (void)b;
  R c = dot3(dx, n);
  R w1 = w_of(ba, rp, n);
  R dl = -c / (w1 + C(1e-6));
  R cm = c < C(0) ? C(1) : C(0);
  R pv[3], q4[4];
  for (int k = 0; k < 3; k++) pv[k] = dl * n[k] * cm;
  half_vqm(ba, rp, pv, qp->rot[a], C(.5), q4);
  for (int k = 0; k < 3; k++) dqa->pos[k] = s->collide_scale * (pv[k] / ba->mass);
  for (int k = 0; k < 4; k++) dqa->rot[k] = s->collide_scale * q4[k];
  /* static friction */
  R qinv[4] = {qp->rot[a][0] * C(1), qp->rot[a][1] * C(-1), qp->rot[a][2] * C(-1),
               qp->rot[a][3] * C(-1)};
  R r1[3], r1o[3], dpv[3], dpt[3];
  rotate(rp, qinv, r1);
  rotate(r1, qprev->rot[a], r1o);
  for (int k = 0; k < 3; k++) dpv[k] = ct->pos[k] - (qprev->pos[a][k] + r1o[k]);
  R dn = dot3(dpv, n);
  for (int k = 0; k < 3; k++) dpt[k] = dpv[k] - dn * n[k];
  R c2 = safe_norm3(s, dpt);
  R n2[3];
  for (int k = 0; k < 3; k++) n2[k] = dpt[k] / (c2 + C(1e-6));
  R w1t = w_of(ba, rp, n2);
  R dlt = -c2 / (w1t + C(0.));
  R sm = FABS(dlt) < FABS(p->friction * dl) ? C(1) : C(0);
  R pt[3];
  for (int k = 0; k < 3; k++) pt[k] = dlt * n2[k] * sm * cm;
  half_vqm(ba, rp, pt, qp->rot[a], C(.5), q4);
  for (int k = 0; k < 3; k++)
    dqa->pos[k] = dqa->pos[k] + s->collide_scale * (pt[k] / ba->mass);
  for (int k = 0; k < 4; k++) dqa->rot[k] = dqa->rot[k] + s->collide_scale * q4[k];
  return dl * cm;
}

/* OneWayCollider._velocity_contact colliders.py:379-442 */
static void oneway_velocity(const osys* s, const opair* p, const oqp* qp,
                            const oqp* qold, const ocontact* ct, R dlambda,
                            R* dvel, R* dang) {
  int a = p->body_a;
  const obody* ba = &s->bodies[a];
  R n[3], r[3], cr[3], rel[3], vt[3], vdir[3];
  for (int k = 0; k < 3; k++) { n[k] = ct->normal[k]; r[k] = ct->pos[k] - qp->pos[a][k]; }
  cross3(qp->ang[a], r, cr);
  for (int k = 0; k < 3; k++) rel[k] = qp->vel[a][k] + cr[k];
  R v_n = dot3(rel, n);
  for (int k = 0; k < 3; k++) vt[k] = rel[k] - n[k] * v_n;
  R vtn = safe_norm3(s, vt);
  for (int k = 0; k < 3; k++) vdir[k] = vt[k] / (C(1e-6) + vtn);
  R lim = p->friction * FABS(dlambda) / (C(2.) * s->dt);
  R mn = lim < vtn ? lim : vtn;
  R dv[3], angw[3], pdyn[3];
  for (int k = 0; k < 3; k++) dv[k] = -vdir[k] * mn;
  cross3(r, vdir, angw);
  R w = (C(1) / ba->mass) + dot3(angw, angw);
  for (int k = 0; k < 3; k++) pdyn[k] = dv[k] / (w + C(1e-6));
  /* restitution */
  R ro[3], cro[3], relo[3];
  for (int k = 0; k < 3; k++) ro[k] = ct->pos[k] - qold->pos[a][k];
  cross3(qold->ang[a], ro, cro);
  for (int k = 0; k < 3; k++) relo[k] = qold->vel[a][k] + cro[k];
  R v_n_old = dot3(relo, n);
  R e = p->elasticity * v_n_old;
  R me = e < C(0) ? e : C(0);
  R dx[3], n2[3];
  for (int k = 0; k < 3; k++) dx[k] = n[k] * (-v_n - me);
  R c = safe_norm3(s, dx);
  for (int k = 0; k < 3; k++) n2[k] = dx[k] / (c + C(1e-6));
  R w1 = w_of(ba, r, n2);
  R dlr = c / (w1 + C(1e-6));
  R stat = ct->pen > C(0) ? C(1) : C(0);
  R sink = v_n_old <= -s->vth ? C(1) : C(0);
  R pv[3], ir[3];
  for (int k = 0; k < 3; k++) pv[k] = (dlr * n2[k] * sink + pdyn[k]) * stat;
  for (int k = 0; k < 3; k++) { dvel[k] = pv[k] / ba->mass; ir[k] = ba->inv_inertia[k] * r[k]; }
  cross3(ir, pv, dang);
}

/* TwoWayCollider._position_contact colliders.py:495-582 */
static R twoway_position(const osys* s, const opair* p, const oqp* qp,
                         const oqp* qprev, const ocontact* ct, dq1* dqa, dq1* dqb) {
  int a = p->body_a, b = p->body_b;
  const obody* ba = &s->bodies[a];
  const obody* bb = &s->bodies[b];
  R rp[3], rc[3], n[3];
  for (int k = 0; k < 3; k++) {
    n[k] = ct->normal[k];
    rp[k] = (ct->pos[k] - ct->normal[k] * ct->pen / C(2.)) - qp->pos[a][k];
    rc[k] = (ct->pos[k] + ct->normal[k] * ct->pen / C(2.)) - qp->pos[b][k];
  }
  R c = -ct->pen;
  R w1 = w_of(ba, rp, n), w2 = w_of(bb, rc, n);
  R dl = -c / (w1 + w2 + C(1e-6));
  R cm = c < C(0) ? C(1) : C(0);
  R pv[3], npv[3], qa4[4], qb4[4];
  for (int k = 0; k < 3; k++) { pv[k] = dl * n[k] * cm; }
  half_vqm(ba, rp, pv, qp->rot[a], C(.5), qa4);
  half_vqm(bb, rc, pv, qp->rot[b], C(-.5), qb4);
  for (int k = 0; k < 3; k++) {
    dqa->pos[k] = s->collide_scale * (pv[k] / ba->mass);
    dqb->pos[k] = s->collide_scale * (-pv[k] / bb->mass);
  }
  for (int k = 0; k < 4; k++) {
    dqa->rot[k] = s->collide_scale * qa4[k];
    dqb->rot[k] = s->collide_scale * qb4[k];
  }
  /* static friction */
  R qai[4] = {qp->rot[a][0] * C(1), qp->rot[a][1] * C(-1), qp->rot[a][2] * C(-1), qp->rot[a][3] * C(-1)};
  R qbi[4] = {qp->rot[b][0] * C(1), qp->rot[b][1] * C(-1), qp->rot[b][2] * C(-1), qp->rot[b][3] * C(-1)};
  R ra[3], rb[3], r1[3], r2[3], r1o[3], r2o[3], dpv[3], dpt[3];
  for (int k = 0; k < 3; k++) { ra[k] = ct->pos[k] - qp->pos[a][k]; rb[k] = ct->pos[k] - qp->pos[b][k]; }
  rotate(ra, qai, r1);
  rotate(rb, qbi, r2);
  rotate(r1, qprev->rot[a], r1o);
  rotate(r2, qprev->rot[b], r2o);
  for (int k = 0; k < 3; k++) {
    R p1bar = qprev->pos[a][k] + r1o[k];
    R p2bar = qprev->pos[b][k] + r2o[k];
    dpv[k] = (ct->pos[k] - p1bar) - (ct->pos[k] - p2bar);
  }
  R dn = dot3(dpv, n);
  for (int k = 0; k < 3; k++) dpt[k] = dpv[k] - dn * n[k];
  R c2 = safe_norm3(s, dpt);
  R n2[3];
  for (int k = 0; k < 3; k++) n2[k] = dpt[k] / (c2 + C(1e-6));
  R w1t = w_of(ba, ra, n2), w2t = w_of(bb, rb, n2);
  R dlt = -c2 / (w1t + w2t);
  R sm = FABS(dlt) < FABS(dl) ? C(1) : C(0);
  R pt[3];
  for (int k = 0; k < 3; k++) { pt[k] = dlt * n2[k] * sm * cm; npv[k] = -pt[k]; }
  half_vqm(ba, ra, pt, qp->rot[a], C(.5), qa4);
  half_vqm(bb, rb, npv, qp->rot[b], C(.5), qb4);
  for (int k = 0; k < 3; k++) {
    dqa->pos[k] = dqa->pos[k] + s->collide_scale * (pt[k] / ba->mass);
    dqb->pos[k] = dqb->pos[k] + s->collide_scale * (-pt[k] / bb->mass);
  }
  for (int k = 0; k < 4; k++) {
    dqa->rot[k] = dqa->rot[k] + s->collide_scale * qa4[k];
    dqb->rot[k] = dqb->rot[k] + s->collide_scale * qb4[k];
  }
  return dl; /* un-masked (colliders.py:580) */
}

/* TwoWayCollider._velocity_contact colliders.py:584-658 */
static void twoway_velocity(const osys* s, const opair* p, const oqp* qp,
                            const oqp* qold, const ocontact* ct, R dlambda,
                            R* dva, R* daa, R* dvb, R* dab) {
  int a = p->body_a, b = p->body_b;
  const obody* ba = &s->bodies[a];
  const obody* bb = &s->bodies[b];
  R n[3], ra[3], rb[3], va[3], vb[3], rel[3], vt[3], vdir[3];
  for (int k = 0; k < 3; k++) { n[k] = ct->normal[k]; ra[k] = ct->pos[k] - qp->pos[a][k]; rb[k] = ct->pos[k] - qp->pos[b][k]; }
  world_velocity(qp, a, ct->pos, va);
  world_velocity(qp, b, ct->pos, vb);
  for (int k = 0; k < 3; k++) rel[k] = va[k] - vb[k];
  R v_n = dot3(rel, n);
  for (int k = 0; k < 3; k++) vt[k] = rel[k] - n[k] * v_n;
  R vtn = safe_norm3(s, vt);
  for (int k = 0; k < 3; k++) vdir[k] = vt[k] / (C(1e-6) + vtn);
  R lim = p->friction * FABS(dlambda) / (C(2.) * s->dt);
  R mn = lim < vtn ? lim : vtn;
  R dv[3], pdyn[3];
  for (int k = 0; k < 3; k++) dv[k] = -vdir[k] * mn;
  R w1 = w_of(ba, ra, vdir), w2 = w_of(bb, rb, vdir);
  for (int k = 0; k < 3; k++) pdyn[k] = dv[k] / (w1 + w2 + C(1e-6));
  R vao[3], vbo[3], relo[3];
  world_velocity(qold, a, ct->pos, vao);
  world_velocity(qold, b, ct->pos, vbo);
  for (int k = 0; k < 3; k++) relo[k] = vao[k] - vbo[k];
  R v_n_old = dot3(relo, n);
  R e = p->elasticity * v_n_old;
  R me = e < C(0) ? e : C(0);
  R dx[3], n2[3], rc[3];
  for (int k = 0; k < 3; k++) {
    dx[k] = n[k] * (-v_n - me);
    rc[k] = (ct->pos[k] + ct->normal[k] * ct->pen) - qp->pos[b][k];
  }
  R c = safe_norm3(s, dx);
  for (int k = 0; k < 3; k++) n2[k] = dx[k] / (c + C(1e-6));
  R w1r = w_of(ba, ra, n2), w2r = w_of(bb, rc, n2);
  R dlr = c / (w1r + w2r + C(1e-6));
  R stat = ct->pen > C(0) ? C(1) : C(0);
  R sink = v_n_old <= C(0.) ? C(1) : C(0);
  R pv[3], npv[3], ia[3], ib[3];
  for (int k = 0; k < 3; k++) {
    pv[k] = (dlr * n2[k] * sink + pdyn[k]) * stat;
    npv[k] = -pv[k];
    dva[k] = pv[k] / ba->mass;
    dvb[k] = -pv[k] / bb->mass;
    ia[k] = ba->inv_inertia[k] * ra[k];
    ib[k] = bb->inv_inertia[k] * rb[k];
  }
  cross3(ia, pv, daa);
  cross3(ib, npv, dab);
}

/* ---- scatter-and-average over a collider group -------------------------- */
static int any_nz3(const R* v) { return v[0] != C(0) || v[1] != C(0) || v[2] != C(0); }

/* Collider.position_apply colliders.py:198-240, summed over groups
 * (system.py:299-303).  contacts/dlambda: K rows in system order. */
static void colliders_position_apply(const osys* s, const oqp* qp, const oqp* qprev,
                                     oq* out, ocontact* contacts, R* dlambda) {
  zero_q(s, out);
  int p0 = 0, k0 = 0;
  while (p0 < s->np) {
    int g = s->pairs[p0].group, p1 = p0;
    while (p1 < s->np && s->pairs[p1].group == g) p1++;
    oq grp;
    R cnt[MAXB];
    zero_q(s, &grp);
    for (int b = 0; b < s->nb; b++) cnt[b] = C(0);
    static _Thread_local dq1 da[MAXK], db[MAXK];
    int two_way = s->pairs[p0].two_way;
    int k = k0;
    for (int pi = p0; pi < p1; pi++) {
      const opair* p = &s->pairs[pi];
      pair_contacts(s, p, qp, contacts + k);
      for (int c = 0; c < p->n_contacts; c++, k++) {
        if (two_way)
          dlambda[k] = twoway_position(s, p, qp, qprev, &contacts[k], &da[k - k0], &db[k - k0]);
        else
          dlambda[k] = oneway_position(s, p, qp, qprev, &contacts[k], &da[k - k0]);
      }
    }
    /* segment_sum over concat(a rows, b rows) */
    k = k0;
    for (int pi = p0; pi < p1; pi++)
      for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
        int b = s->pairs[pi].body_a;
        cnt[b] += any_nz3(da[k - k0].pos) ? C(1) : C(0);
        for (int q = 0; q < 3; q++) grp.pos[b][q] += da[k - k0].pos[q];
        for (int q = 0; q < 4; q++) grp.rot[b][q] += da[k - k0].rot[q];
      }
    if (two_way) {
      k = k0;
      for (int pi = p0; pi < p1; pi++)
        for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
          int b = s->pairs[pi].body_b;
          cnt[b] += any_nz3(db[k - k0].pos) ? C(1) : C(0);
          for (int q = 0; q < 3; q++) grp.pos[b][q] += db[k - k0].pos[q];
          for (int q = 0; q < 4; q++) grp.rot[b][q] += db[k - k0].rot[q];
        }
    }
    for (int b = 0; b < s->nb; b++) {
      R d = C(1e-6) + cnt[b];
      for (int q = 0; q < 3; q++) grp.pos[b][q] = grp.pos[b][q] / d;
      for (int q = 0; q < 4; q++) grp.rot[b][q] = grp.rot[b][q] / d;
    }
    add_q(s, out, &grp);
    k0 = k;
    p0 = p1;
  }
}

/* Collider.velocity_apply colliders.py:155-196, summed over groups
 * (system.py:309-312) */
static void colliders_velocity_apply(const osys* s, const oqp* qp, const oqp* qold,
                                     const ocontact* contacts, const R* dlambda,
                                     op* out) {
  zero_p(s, out);
  int p0 = 0, k0 = 0;
  while (p0 < s->np) {
    int g = s->pairs[p0].group, p1 = p0;
    while (p1 < s->np && s->pairs[p1].group == g) p1++;
    op grp;
    R cnt[MAXB];
    zero_p(s, &grp);
    for (int b = 0; b < s->nb; b++) cnt[b] = C(0);
    static _Thread_local R va[MAXK][3], aa[MAXK][3], vb[MAXK][3], ab[MAXK][3];
    int two_way = s->pairs[p0].two_way;
    int k = k0;
    for (int pi = p0; pi < p1; pi++) {
      const opair* p = &s->pairs[pi];
      for (int c = 0; c < p->n_contacts; c++, k++) {
        if (two_way)
          twoway_velocity(s, p, qp, qold, &contacts[k], dlambda[k], va[k - k0], aa[k - k0], vb[k - k0], ab[k - k0]);
        else
          oneway_velocity(s, p, qp, qold, &contacts[k], dlambda[k], va[k - k0], aa[k - k0]);
      }
    }
    k = k0;
    for (int pi = p0; pi < p1; pi++)
      for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
        int b = s->pairs[pi].body_a;
        cnt[b] += any_nz3(va[k - k0]) ? C(1) : C(0);
        for (int q = 0; q < 3; q++) { grp.vel[b][q] += va[k - k0][q]; grp.ang[b][q] += aa[k - k0][q]; }
      }
    if (two_way) {
      k = k0;
      for (int pi = p0; pi < p1; pi++)
        for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
          int b = s->pairs[pi].body_b;
          cnt[b] += any_nz3(vb[k - k0]) ? C(1) : C(0);
          for (int q = 0; q < 3; q++) { grp.vel[b][q] += vb[k - k0][q]; grp.ang[b][q] += ab[k - k0][q]; }
        }
    }
    for (int b = 0; b < s->nb; b++) {
      R d = C(1e-6) + cnt[b];
      for (int q = 0; q < 3; q++) { grp.vel[b][q] = grp.vel[b][q] / d; grp.ang[b][q] = grp.ang[b][q] / d; }
    }
    add_p(s, out, &grp);
    k0 = k;
    p0 = p1;
  }
}

/* ---- legacy_spring contacts: colliders.py:116-153, 267-304, 449-493 ------ */
/* Body.impulse bodies.py:46-59 */
static void body_impulse(const obody* bd, const R* bpos, const R* imp, const R* pos,
                         R* dvel, R* dang) {
  R r[3], cr[3];
  for (int k = 0; k < 3; k++) { dvel[k] = imp[k] / bd->mass; r[k] = pos[k] - bpos[k]; }
  cross3(r, imp, cr);
  for (int k = 0; k < 3; k++) dang[k] = bd->inv_inertia[k] * cr[k];
}

static void spring_contact(const osys* s, const opair* p, const oqp* qp,
                           const ocontact* ct, R* va, R* aa, R* vb, R* ab) {
  int a = p->body_a, b = p->body_b;
  const obody* ba = &s->bodies[a];
  const obody* bb = &s->bodies[b];
  R ra[3], rb[3], t1[3], t2[3], c1[3], c2[3], sum[3];
  for (int k = 0; k < 3; k++) { ra[k] = ct->pos[k] - qp->pos[a][k]; rb[k] = ct->pos[k] - qp->pos[b][k]; }
  R baum = s->erp * ct->pen;
  R nv = dot3(ct->normal, ct->vel);
  cross3(ra, ct->normal, t1);
  for (int k = 0; k < 3; k++) t1[k] = ba->inv_inertia[k] * t1[k];
  cross3(t1, ra, c1);
  R ang, denom;
  if (p->two_way) {
    cross3(rb, ct->normal, t2);
    for (int k = 0; k < 3; k++) t2[k] = bb->inv_inertia[k] * t2[k];
    cross3(t2, rb, c2);
    for (int k = 0; k < 3; k++) sum[k] = c1[k] + c2[k];
    ang = dot3(ct->normal, sum);
    denom = (C(1.) / ba->mass) + (C(1.) / bb->mass) + ang;
  } else {
    ang = dot3(ct->normal, c1);
    denom = (C(1.) / ba->mass) + ang;
  }
  R imp = (C(-1.) * (C(1.) + p->elasticity) * nv + baum) / denom;
  R in[3], nin[3], dvn[3], dan[3], dvnb[3], danb[3];
  for (int k = 0; k < 3; k++) { in[k] = imp * ct->normal[k]; nin[k] = -imp * ct->normal[k]; }
  body_impulse(ba, qp->pos[a], in, ct->pos, dvn, dan);
  R vd[3], dir[3], idv[3], nidv[3], dvd[3], dad[3], dvdb[3], dadb[3];
  for (int k = 0; k < 3; k++) vd[k] = ct->vel[k] - nv * ct->normal[k];
  R nvd = safe_norm3(s, vd);
  R impd = nvd / denom;
  R cap = p->friction * imp;
  impd = impd < cap ? impd : cap; /* jp.minimum */
  for (int k = 0; k < 3; k++) { dir[k] = vd[k] / (C(1e-6) + nvd); idv[k] = -impd * dir[k]; nidv[k] = impd * dir[k]; }
  body_impulse(ba, qp->pos[a], idv, ct->pos, dvd, dad);
  R apn = (ct->pen > C(0.) && nv < C(0) && imp > C(0.)) ? C(1) : C(0);
  R apd = apn * (nvd > C(0.01) ? C(1) : C(0));
  for (int k = 0; k < 3; k++) {
    va[k] = dvn[k] * apn + dvd[k] * apd;
    aa[k] = dan[k] * apn + dad[k] * apd;
  }
  if (p->two_way) {
    body_impulse(bb, qp->pos[b], nin, ct->pos, dvnb, danb);
    body_impulse(bb, qp->pos[b], nidv, ct->pos, dvdb, dadb);
    for (int k = 0; k < 3; k++) {
      vb[k] = dvnb[k] * apn + dvdb[k] * apd;
      ab[k] = danb[k] * apn + dadb[k] * apd;
    }
  }
}

/* Collider.apply colliders.py:116-153 summed over groups (system.py:362-363) */
static void colliders_apply_spring(const osys* s, const oqp* qp, op* out,
                                   ocontact* contacts) {
  zero_p(s, out);
  int p0 = 0, k0 = 0;
  while (p0 < s->np) {
    int g = s->pairs[p0].group, p1 = p0;
    while (p1 < s->np && s->pairs[p1].group == g) p1++;
    op grp;
    R cnt[MAXB];
    zero_p(s, &grp);
    for (int b = 0; b < s->nb; b++) cnt[b] = C(0);
    static _Thread_local R va[MAXK][3], aa[MAXK][3], vb[MAXK][3], ab[MAXK][3];
    int two_way = s->pairs[p0].two_way;
    int k = k0;
    for (int pi = p0; pi < p1; pi++) {
      const opair* p = &s->pairs[pi];
      pair_contacts(s, p, qp, contacts + k);
      for (int c = 0; c < p->n_contacts; c++, k++)
        spring_contact(s, p, qp, &contacts[k], va[k - k0], aa[k - k0], vb[k - k0], ab[k - k0]);
    }
    k = k0;
    for (int pi = p0; pi < p1; pi++)
      for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
        int b = s->pairs[pi].body_a;
        cnt[b] += any_nz3(va[k - k0]) ? C(1) : C(0);
        for (int q = 0; q < 3; q++) { grp.vel[b][q] += va[k - k0][q]; grp.ang[b][q] += aa[k - k0][q]; }
      }
    if (two_way) {
      k = k0;
      for (int pi = p0; pi < p1; pi++)
        for (int c = 0; c < s->pairs[pi].n_contacts; c++, k++) {
          int b = s->pairs[pi].body_b;
          cnt[b] += any_nz3(vb[k - k0]) ? C(1) : C(0);
          for (int q = 0; q < 3; q++) { grp.vel[b][q] += vb[k - k0][q]; grp.ang[b][q] += ab[k - k0][q]; }
        }
    }
    for (int b = 0; b < s->nb; b++) {
      R d = C(1e-8) + cnt[b];
      for (int q = 0; q < 3; q++) { grp.vel[b][q] = grp.vel[b][q] / d; grp.ang[b][q] = grp.ang[b][q] / d; }
    }
    add_p(s, out, &grp);
    k0 = k;
    p0 = p1;
  }
}

/* ---- legacy_spring joints: spring_joints.py:89-303 ------------------------ */
static void to_world(const oqp* qp, int b, const R* off, R* wpos, R* wvel) {
  /* QP.to_world base.py:112-125 */
  R ro[3], cr[3];
  rotate(off, qp->rot[b], ro);
  cross3(qp->ang[b], ro, cr);
  for (int k = 0; k < 3; k++) { wpos[k] = qp->pos[b][k] + ro[k]; wvel[k] = qp->vel[b][k] + cr[k]; }
}

static void spring_joint_apply(const osys* s, const ojoint* j, const oqp* qp,
                               R* vp, R* ap, R* vc, R* ac) {
  int bpi = j->body_p, bci = j->body_c;
  const obody* bp = &s->bodies[bpi];
  const obody* bc = &s->bodies[bci];
  R pp[3], vpw[3], pc[3], vcw[3], imp[3], nimp[3];
  to_world(qp, bpi, j->off_p, pp, vpw);
  to_world(qp, bci, j->off_c, pc, vcw);
  for (int k = 0; k < 3; k++) {
    imp[k] = (pp[k] - pc[k]) * j->stiffness + j->spring_damping * (vpw[k] - vcw[k]);
    nimp[k] = -imp[k];
  }
  body_impulse(bp, qp->pos[bpi], nimp, pp, vp, ap);
  body_impulse(bc, qp->pos[bci], imp, pc, vc, ac);
  R axis[3][3], angle[3], tq[3];
  joint_axis_angle(s, j, qp->rot[bpi], qp->rot[bci], axis, angle);
  int dof = joint_dof(j);
  R dang[3];
  for (int d = 0; d < dof; d++) {
    R v = C(0);
    if (angle[d] < j->limit[2 * d]) v = j->limit[2 * d] - angle[d];
    if (angle[d] > j->limit[2 * d + 1]) v = j->limit[2 * d + 1] - angle[d];
    dang[d] = v;
  }
  if (j->kind == J_SREV) { /* spring_joints.py:127-148 */
    R axc[3], cr[3];
    rotate(j->axis_c + 0, qp->rot[bci], axc);
    cross3(axis[0], axc, cr);
    for (int k = 0; k < 3; k++) {
      tq[k] = j->stiffness * cr[k];
      tq[k] = tq[k] - j->limit_strength * axis[0][k] * dang[0];
    }
  } else if (j->kind == J_SUNI) { /* spring_joints.py:178-210 */
    R d21 = dot3(axis[1], axis[0]), proj[3], cr[3];
    for (int k = 0; k < 3; k++) proj[k] = axis[1][k] - d21 * axis[0][k];
    R np_ = safe_norm3(s, proj);
    for (int k = 0; k < 3; k++) proj[k] = proj[k] / np_;
    cross3(proj, axis[1], cr);
    for (int k = 0; k < 3; k++) {
      tq[k] = (j->limit_strength / C(5.)) * cr[k];
      R sm = axis[0][k] * dang[0] + axis[1][k] * dang[1];
      tq[k] = tq[k] - j->limit_strength * sm;
    }
  } else { /* spring Spherical spring_joints.py:249-274 */
    for (int k = 0; k < 3; k++) {
      R sm = axis[0][k] * dang[0] + axis[1][k] * dang[1] + axis[2][k] * dang[2];
      tq[k] = -j->limit_strength * sm;
    }
  }
  for (int k = 0; k < 3; k++) {
    tq[k] = tq[k] - j->angular_damping * (qp->ang[bpi][k] - qp->ang[bci][k]);
    ap[k] = ap[k] + bp->inv_inertia[k] * tq[k];
    ac[k] = ac[k] + -bc->inv_inertia[k] * tq[k];
  }
}

static void joints_apply_spring(const osys* s, const oqp* qp, op* out) {
  zero_p(s, out);
  int i0 = 0;
  while (i0 < s->nj) {
    int g = s->joints[i0].group, i1 = i0;
    while (i1 < s->nj && s->joints[i1].group == g) i1++;
    if (s->joints[i0].kind >= J_SREV) {
      op grp;
      zero_p(s, &grp);
      R vp[64][3], ap[64][3], vc[64][3], ac[64][3];
      for (int i = i0; i < i1; i++)
        spring_joint_apply(s, &s->joints[i], qp, vp[i - i0], ap[i - i0], vc[i - i0], ac[i - i0]);
      for (int i = i0; i < i1; i++) {
        int b = s->joints[i].body_p;
        for (int k = 0; k < 3; k++) { grp.vel[b][k] += vp[i - i0][k]; grp.ang[b][k] += ap[i - i0][k]; }
      }
      for (int i = i0; i < i1; i++) {
        int b = s->joints[i].body_c;
        for (int k = 0; k < 3; k++) { grp.vel[b][k] += vc[i - i0][k]; grp.ang[b][k] += ac[i - i0][k]; }
      }
      add_p(s, out, &grp);
    }
    i0 = i1;
  }
}

/* ---- the step: system.py:244-377 --------------------------------------- */
typedef struct {
  op contact, joint, actuator;
  ocontact contacts[MAXK];
  int have_contacts;
} oinfo;

static int num_contacts(const osys* s) {
  int k = 0;
  for (int i = 0; i < s->np; i++) k += s->pairs[i].n_contacts;
  return k;
}

/* _pbd_step system.py:254-325 */
static void pbd_step(const osys* s, oqp* qp, const R* act, oinfo* info) {
  static _Thread_local oqp qprev, qrb;
  static _Thread_local op dpa, dpf, dpj, dpc;
  static _Thread_local oq dq;
  static _Thread_local R dlambda[MAXK];
  zero_p(s, &info->contact);
  zero_p(s, &info->joint);
  zero_p(s, &info->actuator);
  info->have_contacts = 0;
  for (int it = 0; it < s->substeps / 2; it++) {
    /* half A, no collisions: system.py:262-281 */
    qprev = *qp;
    actuators_apply(s, qp, act, &dpa);
    forces_apply(s, act, &dpf);
    joints_damp(s, qp, &dpj);
    add_p(s, &dpa, &dpf);
    add_p(s, &dpa, &dpj);
    update_acc(s, qp, &dpa);
    kinetic(s, qp);
    joints_apply_pbd(s, qp, &dq);
    update_pos(s, qp, &dq);
    velocity_projection(s, qp, &qprev);
    /* half B, with collisions: system.py:283-313 */
    qprev = *qp;
    actuators_apply(s, qp, act, &dpa);
    forces_apply(s, act, &dpf);
    joints_damp(s, qp, &dpj);
    {
      op acc = dpa;
      add_p(s, &acc, &dpf);
      add_p(s, &acc, &dpj);
      update_acc(s, qp, &acc);
    }
    kinetic(s, qp);
    joints_apply_pbd(s, qp, &dq);
    update_pos(s, qp, &dq);
    colliders_position_apply(s, qp, &qprev, &dq, info->contacts, dlambda);
    update_pos(s, qp, &dq);
    qrb = *qp;
    velocity_projection(s, qp, &qprev);
    colliders_velocity_apply(s, qp, &qrb, info->contacts, dlambda, &dpc);
    update_vel(s, qp, &dpc);
    add_p(s, &info->contact, &dpc);
    add_p(s, &info->actuator, &dpa); /* half-B dp_a only, system.py:287,315 */
    info->have_contacts = 1;
  }
}

/* _spring_step system.py:342-377 */
static void spring_step(const osys* s, oqp* qp, const R* act, oinfo* info) {
  static _Thread_local op dpa, dpf, dpj, dpc;
  zero_p(s, &info->contact);
  zero_p(s, &info->joint);
  zero_p(s, &info->actuator);
  info->have_contacts = 0;
  for (int it = 0; it < s->substeps; it++) {
    kinetic(s, qp);
    joints_apply_spring(s, qp, &dpj);
    actuators_apply(s, qp, act, &dpa);
    forces_apply(s, act, &dpf);
    {
      op acc = dpj;
      add_p(s, &acc, &dpa);
      add_p(s, &acc, &dpf);
      update_acc(s, qp, &acc);
    }
    colliders_apply_spring(s, qp, &dpc, info->contacts);
    update_vel(s, qp, &dpc);
    add_p(s, &info->contact, &dpc);
    add_p(s, &info->joint, &dpj);
    add_p(s, &info->actuator, &dpa);
    info->have_contacts = 1;
  }
}

/* ---- exported entry points ---------------------------------------------- */
typedef struct {
  const R *pos, *rot, *vel, *ang, *act;
  R *pos_o, *rot_o, *vel_o, *ang_o;
  R *contact_vel, *contact_ang, *joint_vel, *joint_ang, *actuator_vel,
      *actuator_ang, *contact_pos, *contact_normal, *contact_penetration;
} oio;

int FN(bxo_num_contacts)(const osys* s) { return num_contacts(s); }

static void step_one(const osys* s, const oio* io, int64_t e) {
  static _Thread_local oqp qp;
  static _Thread_local oinfo info;
  int n = s->nb, K = num_contacts(s);
  const R* pos = io->pos + e * n * 3;
  const R* rot = io->rot + e * n * 4;
  const R* vel = io->vel + e * n * 3;
  const R* ang = io->ang + e * n * 3;
  for (int b = 0; b < n; b++) {
    for (int k = 0; k < 3; k++) {
      qp.pos[b][k] = pos[b * 3 + k];
      qp.vel[b][k] = vel[b * 3 + k];
      qp.ang[b][k] = ang[b * 3 + k];
    }
    for (int k = 0; k < 4; k++) qp.rot[b][k] = rot[b * 4 + k];
  }
  const R* act = io->act + e * s->act_size;
  if (s->mode == MODE_PBD) pbd_step(s, &qp, act, &info);
  else spring_step(s, &qp, act, &info);
  for (int b = 0; b < n; b++) {
    for (int k = 0; k < 3; k++) {
      io->pos_o[(e * n + b) * 3 + k] = qp.pos[b][k];
      io->vel_o[(e * n + b) * 3 + k] = qp.vel[b][k];
      io->ang_o[(e * n + b) * 3 + k] = qp.ang[b][k];
      if (io->contact_vel) io->contact_vel[(e * n + b) * 3 + k] = info.contact.vel[b][k];
      if (io->contact_ang) io->contact_ang[(e * n + b) * 3 + k] = info.contact.ang[b][k];
      if (io->joint_vel) io->joint_vel[(e * n + b) * 3 + k] = info.joint.vel[b][k];
      if (io->joint_ang) io->joint_ang[(e * n + b) * 3 + k] = info.joint.ang[b][k];
      if (io->actuator_vel) io->actuator_vel[(e * n + b) * 3 + k] = info.actuator.vel[b][k];
      if (io->actuator_ang) io->actuator_ang[(e * n + b) * 3 + k] = info.actuator.ang[b][k];
    }
    for (int k = 0; k < 4; k++) io->rot_o[(e * n + b) * 4 + k] = qp.rot[b][k];
  }
  for (int c = 0; c < K; c++) {
    /* zero_info: pos/normal zeros, penetration -1 (system.py:78-84) */
    for (int k = 0; k < 3; k++) {
      if (io->contact_pos)
        io->contact_pos[(e * K + c) * 3 + k] = info.have_contacts ? info.contacts[c].pos[k] : C(0);
      if (io->contact_normal)
        io->contact_normal[(e * K + c) * 3 + k] = info.have_contacts ? info.contacts[c].normal[k] : C(0);
    }
    if (io->contact_penetration)
      io->contact_penetration[e * K + c] = info.have_contacts ? info.contacts[c].pen : C(-1);
  }
}

/* One System.step for `num_envs` envs; layout as include/brax_b200.h.
 * OpenMP over envs when built with -fopenmp (num_threads <= 0: all cores). */
int FN(bxo_step)(const osys* s, int64_t num_envs, const oio* io, int num_threads) {
  if (s->nb > MAXB || num_contacts(s) > MAXK || s->nj > 64 || s->na > 64) return 1;
  (void)num_threads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(num_threads > 0 ? num_threads : omp_get_max_threads_wrap())
#endif
  for (int64_t e = 0; e < num_envs; e++) step_one(s, io, e);
  return 0;
}

/* math.rotate (math.py:25-40) for the reference's known-answer test math_test.py:25-37. */
void FN(bxo_rotate)(const R* v, const R* q, R* out) { rotate(v, q, out); }

/* contact_fn of pair `pair_index` evaluated on one env's QP (unbatched
 * [N,3|4] arrays); out: n_contacts rows of pos[3] vel[3] normal[3] pen.
 * For the reference's collider known-answer tests (colliders_test.py:54-135). */
int FN(bxo_pair_contacts)(const osys* s, int pair_index, const R* pos, const R* rot,
                          const R* vel, const R* ang, R* out10) {
  static _Thread_local oqp qp;
  ocontact ct[12];
  if (pair_index < 0 || pair_index >= s->np) return 1;
  for (int b = 0; b < s->nb; b++) {
    for (int k = 0; k < 3; k++) {
      qp.pos[b][k] = pos[b * 3 + k];
      qp.vel[b][k] = vel[b * 3 + k];
      qp.ang[b][k] = ang[b * 3 + k];
    }
    for (int k = 0; k < 4; k++) qp.rot[b][k] = rot[b * 4 + k];
  }
  const opair* p = &s->pairs[pair_index];
  pair_contacts(s, p, &qp, ct);
  for (int c = 0; c < p->n_contacts; c++) {
    for (int k = 0; k < 3; k++) {
      out10[c * 10 + k] = ct[c].pos[k];
      out10[c * 10 + 3 + k] = ct[c].vel[k];
      out10[c * 10 + 6 + k] = ct[c].normal[k];
    }
    out10[c * 10 + 9] = ct[c].pen;
  }
  return 0;
}

/* Joint.angle_vel (joints.py:197-226) for every joint, BEFORE the free_dofs
 * selection of :219-225: angle_out / vel_out are [B, nj, 3], the slots beyond a
 * joint's dof stay 0.  vel_d = dot(ang_p - ang_c, axis_d) (:212). */
int FN(bxo_angle_vel)(const osys* s, int64_t num_envs, const R* rot, const R* ang,
                      R* angle_out, R* vel_out) {
  for (int64_t e = 0; e < num_envs; e++) {
    const R* re = rot + e * s->nb * 4;
    const R* ae = ang + e * s->nb * 3;
    for (int i = 0; i < s->nj; i++) {
      const ojoint* j = &s->joints[i];
      R axis[3][3] = {{0}}, angle[3] = {0, 0, 0};
      joint_axis_angle(s, j, re + 4 * j->body_p, re + 4 * j->body_c, axis, angle);
      const int dof = joint_dof(j);
      R* ao = angle_out + (e * s->nj + i) * 3;
      R* vo = vel_out + (e * s->nj + i) * 3;
      for (int d = 0; d < 3; d++) {
        R dw[3];
        for (int k = 0; k < 3; k++) dw[k] = ae[3 * j->body_p + k] - ae[3 * j->body_c + k];
        ao[d] = d < dof ? angle[d] : C(0);
        vo[d] = d < dof ? dot3(dw, axis[d]) : C(0);
      }
    }
  }
  return 0;
}

/* System.info(qp).contact and the contact data (system.py:327-340 `_pbd_info`, :378-390 `_spring_info`):
 * in BOTH dynamics modes the summed `Collider.apply(qp)` of all colliders -- the spring-style
 * contact impulse (colliders.py:116-153) evaluated on the state, with the baumgarte_erp of the
 * config -- plus contact_pos / normal / penetration of every contact row.  This is the Info the
 * reference's Env.reset hands to _get_obs (envs/ant.py:204).  Arrays are batched [B, ...]. */
int FN(bxo_info_contact)(const osys* s, int64_t num_envs, const R* pos, const R* rot, const R* vel,
                         const R* ang, R* cvel, R* cang, R* cpos, R* cnormal, R* cpen) {
  const int nk = num_contacts(s);
  for (int64_t e = 0; e < num_envs; e++) {
    static _Thread_local oqp qp;
    static _Thread_local op dp;
    static _Thread_local ocontact ct[MAXK];
    for (int b = 0; b < s->nb; b++) {
      for (int k = 0; k < 3; k++) {
        qp.pos[b][k] = pos[(e * s->nb + b) * 3 + k];
        qp.vel[b][k] = vel[(e * s->nb + b) * 3 + k];
        qp.ang[b][k] = ang[(e * s->nb + b) * 3 + k];
      }
      for (int k = 0; k < 4; k++) qp.rot[b][k] = rot[(e * s->nb + b) * 4 + k];
    }
    colliders_apply_spring(s, &qp, &dp, ct);
    for (int b = 0; b < s->nb; b++)
      for (int k = 0; k < 3; k++) {
        cvel[(e * s->nb + b) * 3 + k] = dp.vel[b][k];
        cang[(e * s->nb + b) * 3 + k] = dp.ang[b][k];
      }
    for (int c = 0; c < nk; c++) {
      for (int k = 0; k < 3; k++) {
        cpos[(e * nk + c) * 3 + k] = ct[c].pos[k];
        cnormal[(e * nk + c) * 3 + k] = ct[c].normal[k];
      }
      cpen[e * nk + c] = ct[c].pen;
    }
  }
  return 0;
}

#ifdef BXO_COUNT
/* C-linkage entry of the counting build (oracle/count_flops.py) */
extern "C" int bxo_step_count(const osys* s, int64_t num_envs, const oio* io) {
  return FN(bxo_step)(s, num_envs, io, 1);
}
#endif
