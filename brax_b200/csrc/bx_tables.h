// bx_tables.h -- flat device-side "program" of one system.
//
// bx_system_create() (bx_api.cu) flattens a bx_system_desc into one int32 and
// one float32 array (uploaded once); the step kernel copies both into shared
// memory at CTA start.  Every index in here is static per system, so inside the
// kernel all table reads are warp-uniform (one warp = 32 environments working
// on the same joint / body / contact row).
//
// Reference anchors for what each record holds: joints.py:38-77,
// spring_joints.py:42-87, actuators.py:31-46, forces.py:29-41,
// geometry.py:86-93,123-136,243-288, bodies.py:39-43, integrators.py:38-48.
#ifndef BX_TABLES_H_
#define BX_TABLES_H_

#include <stdint.h>

// ---- float records (strides in floats) ------------------------------------
enum {
  // body
  BF_INV_MASS = 0,  // 1/mass (float32 division on the host, as jnp does)
  BF_MASS = 1,
  BF_INV_INERTIA = 2,  // [3]
  BF_POS_MASK = 5,     // [3]
  BF_ROT_MASK = 8,     // [3]
  BF_STRIDE = 11,

  // joint
  JF_OFF_P = 0,    // [3]
  JF_OFF_C = 3,    // [3]
  JF_AXIS_P = 6,   // [9] row-major
  JF_AXIS_C = 15,  // [9]
  JF_LIMIT = 24,   // [3][2]
  JF_ANG_DAMP = 30,
  JF_SCALE_POS = 31,
  JF_SCALE_ANG = 32,
  JF_STIFFNESS = 33,
  JF_SPRING_DAMP = 34,
  JF_LIMIT_STRENGTH = 35,
  JF_STRIDE = 37,

  // contact row (one per contact, pairs expanded)
  CF_FRICTION = 0,
  CF_ELASTICITY = 1,
  CF_RADIUS_A = 2,
  CF_RADIUS_B = 3,
  CF_PTS = 4,  // capsule_plane / box_plane: [3]; capsule_capsule: [4][3]
  CF_STRIDE = 17,
};

// ---- int records (strides in ints) -----------------------------------------
enum {
  // body
  BI_CSLOT = 0,     // slot in the collider-body regions (PRE, INFO_C) or -1
  BI_ASLOT = 1,     // slot in the actuated-body region (INFO_A) or -1
  BI_JADJ_OFF = 2,  // joint adjacency: entries (j << 1 | side), side 1 = child
  BI_JADJ_N = 3,
  BI_AADJ_OFF = 4,  // actuator adjacency: (a << 2 | negate << 1 | side)
  BI_AADJ_N = 5,
  BI_FADJ_OFF = 6,  // force adjacency: force index
  BI_FADJ_N = 7,
  BI_CADJ_OFF = 8,  // contact adjacency: (row << 4 | group << 1 | side)
  BI_CADJ_N = 9,
  BI_STRIDE = 10,

  // joint
  JI_KIND = 0,
  JI_BODY_P = 1,
  JI_BODY_C = 2,
  JI_STRIDE = 3,

  // actuator
  AI_KIND = 0,
  AI_JOINT = 1,
  AI_IDX = 2,  // [3], -1 = masked
  AI_STRIDE = 5,

  // force
  FI_KIND = 0,
  FI_BODY = 1,
  FI_IDX = 2,  // [3]
  FI_STRIDE = 5,

  // contact row
  CI_KIND = 0,
  CI_TWO_WAY = 1,
  CI_BODY_A = 2,
  CI_BODY_B = 3,
  CI_DELTA_OFF = 4,  // row offset of this contact's delta slot in region JS
  CI_STRIDE = 5,
};

// Header: lives in the kernel parameter (constant bank), not in shared memory.
struct BxHeader {
  int32_t mode;      // BX_MODE_*
  int32_t iters;     // pbd: substeps/2 ; legacy_spring: substeps
  int32_t nb, nj, na, nf, nk;  // bodies, joints, actuators, forces, contact rows
  int32_t act_size;
  int32_t ncb;  // bodies that appear in a collider (slots for PRE/INFO_C)
  int32_t nab;  // bodies that appear in an actuated joint (slots for INFO_A)
  // table sizes / offsets
  int32_t n_ints, n_floats;
  int32_t io_body, io_joint, io_act, io_force, io_contact, io_adj;  // int offs
  int32_t fo_body, fo_joint, fo_act, fo_force, fo_contact;          // float offs
  // state rows (row index; each row is 33 floats = 32 envs + 1 pad)
  int32_t r_pos0, r_pos1, r_rot0, r_rot1, r_vel, r_ang;
  int32_t r_act, r_js, r_ct, r_pre, r_ic, r_ia, r_ij;
  int32_t n_rows;
  // output maps (offsets into the int table; entry = state row or -1 for 0):
  // contact_vel[3N] contact_ang[3N] actuator_ang[3N] joint_vel[3N] joint_ang[3N]
  // contact_pos[3K] contact_normal[3K] contact_penetration[K]
  int32_t m_cv, m_ca, m_aa, m_jv, m_ja, m_cp, m_cn, m_pen;
  // scalars
  float dt, inv_dt, half_dt;
  float gravity[3];
  float vel_damp, ang_damp;
  float erp, collide_scale, vth;
  float inv_2h;  // 1 / (2*dt_sub), colliders.py:399
};

#define BX_ROW_STRIDE 33  // 32 env slots + 1 pad word (bank-conflict-free I/O)
#define BX_ENVS_PER_CTA 32

#endif  // BX_TABLES_H_
