// bx_tables.h -- flat device-side "program" of one system.
//
// bx_system_create() (bx_api.cu) flattens a bx_system_desc into one float4
// table and one int32 table (uploaded once); the step kernel copies both into
// shared memory at CTA start.  Every index in here is static per system, so
// inside the kernel all table reads are warp-uniform 128-bit shared-memory
// broadcasts (one warp = 32 environments working on the same joint / body /
// contact row).
//
// Reference anchors for what each record holds: joints.py:38-77,
// spring_joints.py:42-87, actuators.py:31-46, forces.py:29-41,
// geometry.py:86-93,123-136,243-288, bodies.py:39-43, integrators.py:38-48.
#ifndef BX_TABLES_H_
#define BX_TABLES_H_

#include <stdint.h>

// ---- float4 records (strides in quads = float4) --------------------------
enum {
  // body: 4 quads
  BQ_MASS = 0,         // {1/mass (float32 division on the host, as jnp does), mass, 0, 0}
  BQ_INV_INERTIA = 1,  // {ix, iy, iz, 0}
  BQ_POS_MASK = 2,     // {x, y, z, 0}
  BQ_ROT_MASK = 3,     // {x, y, z, 0}
  BQ_STRIDE = 4,

  // joint: 10 quads
  JQ_OFF_P = 0,   // {off_p, angular_damping}
  JQ_OFF_C = 1,   // {off_c, scale_pos}
  JQ_AXIS_P = 2,  // rows 0..2: {axis_p[r], w}; w = scale_ang, stiffness, spring_damping
  JQ_AXIS_C = 5,  // rows 0..2: {axis_c[r], w}; w = limit_strength, 0, 0
  JQ_LIMIT = 8,   // {lo0, hi0, lo1, hi1}, {lo2, hi2, 0, 0}
  JQ_LIMCS = 10,  // per dof d: {cos lo_d, sin lo_d, cos hi_d, sin hi_d} (a limit at or beyond +-pi: (-1, 0))
  JQ_STRIDE = 13,
  JQ_STRIDE_REVOLUTE = 10,  // systems with revolute joints only: quad 9 = the dof-0 limit directions

  // actuator / force: 1 quad {strength, 0, 0, 0}
  // contact row (one per contact, pairs expanded): 5 quads
  CQ_MAT = 0,  // {friction, elasticity, radius_a, radius_b}
  CQ_PTS = 1,  // capsule_plane / box_plane: pts[0]; capsule_capsule: pts[0..3]
  CQ_STRIDE = 5,
};

// ---- int records (strides in ints; every record is 16-byte aligned) --------
// "Slots" are quad rows relative to region JS.  Every producer (actuator, joint,
// contact row) writes its per-body contribution straight into the consuming
// body's slot LIST; the lists are contiguous and laid out in the reference's
// summation order (joints.py:93-98, actuators.py:67-70, colliders.py:222-228),
// so a body sums its list with independent, address-known loads.
enum {
  // body: 4 x int4
  BI_CSLOT = 0,   // slot in the collider-body regions (PRE, INFO_C) or -1
  BI_ASLOT = 1,   // slot in the actuated-body region (INFO_A) or -1
  BI_A_BASE = 2,  // actuator list: a_n quads (signed I^-1 tau)
  BI_A_N = 3,
  BI_D_BASE = 4,  // joint-damping list: d_n quads (pbd)
  BI_D_N = 5,
  BI_J_BASE = 6,  // joint list: j_n entries of 2 quads (pbd: dpos, drot; spring: dvel, dang)
  BI_J_N = 7,
  BI_F_OFF = 8,   // force adjacency (offset into the adj table), force indices
  BI_F_N = 9,
  BI_CG_OFF = 10,  // contact groups of this body: cg_n pairs (base, count) in the adj
  BI_CG_N = 11,    // table; each entry = 2 quads (linear part, second part)
  BI_ROW_OFF = 12,  // contact items whose body_a is this body (adj table), fused path
  BI_ROW_N = 13,
  BI_STRIDE = 16,

  // joint: 2 x int4 {kind, body_p, body_c, 0} {dslot_p, dslot_c, jslot_p, jslot_c}
  JI_KIND = 0,
  JI_BODY_P = 1,
  JI_BODY_C = 2,
  JI_DSLOT_P = 4,
  JI_DSLOT_C = 5,
  JI_JSLOT_P = 6,
  JI_JSLOT_C = 7,
  JI_STRIDE = 8,

  // actuator: 2 x int4 {kind, joint, idx0, idx1} {idx2, slot_p, slot_c, 0}
  AI_KIND = 0,
  AI_JOINT = 1,
  AI_IDX = 2,  // [3], -1 = masked
  AI_SLOT_P = 5,
  AI_SLOT_C = 6,
  AI_STRIDE = 8,

  // force: 2 x int4
  FI_KIND = 0,
  FI_BODY = 1,
  FI_IDX = 2,  // [3]
  FI_STRIDE = 8,

  // contact row: 2 x int4 {kind, two_way, body_a, body_b} {slot_a, slot_b, normal row, triangle quad offset}.
  // slot_a / slot_b are those of the row's ITEM (1 or 2 consecutive rows of one pair
  // that share one slot-list entry per side); `pair` is the row of the pair's normal.
  CI_KIND = 0,
  CI_TWO_WAY = 1,
  CI_BODY_A = 2,
  CI_BODY_B = 3,
  CI_SLOT_A = 4,
  CI_SLOT_B = 5,
  CI_PAIR = 6,   // row of the contact normal (per pair; per row for capsule_box)
  CI_TRI = 7,    // capsule_box: quad offset of the row's triangle in the fo_tri table
  CI_STRIDE = 8,
};

// Global <-> shared I/O segment: global array g[B][nper]; element (e, r) lives at
// shared float index map[r] + 4 * e (map[r] < 0: the element is a constant).
#define BX_SEG_IN 5    // pos rot vel ang act
#define BX_SEG_OUT_STEP 15  // pos rot vel ang + 6 Info P + contact pos/normal/pen + joint angle/velocity
#define BX_SEG_OUT 19       // ... + env epilogue: obs, reward, done, metrics (written every step of a rollout)
#define BX_SEG_ENV BX_SEG_OUT_STEP
struct BxSeg {
  int32_t nper;   // floats per environment
  int32_t map;    // offset of map[nper] in the int table
  uint32_t magic; // ceil(2^32 / nper): e = umulhi(idx, magic) for idx < 2^20
  float fill;     // value of elements with map[r] < 0
};

// Header: lives in the kernel parameter (constant bank), not in shared memory.
struct BxHeader {
  int32_t mode;      // BX_MODE_*
  int32_t iters;     // pbd: substeps/2 ; legacy_spring: substeps
  int32_t nb, nj, na, nf, nk;  // bodies, joints, actuators, forces, contact rows
  int32_t nci;       // contact items (1 or 2 rows each); io_citem: {first row, rows} per item
  int32_t act_size;
  int32_t ncb;  // bodies that appear in a collider (slots for PRE/INFO_C)
  int32_t nab;  // bodies that appear in an actuated joint (slots for INFO_A)
  // table sizes / offsets
  int32_t n_ints, n_fquads;
  int32_t io_body, io_joint, io_act, io_force, io_contact, io_adj;  // int offs
  int32_t io_forder;  // body order of the fused contact section (contact bodies first)
  int32_t io_citem;
  int32_t fo_body, fo_joint, fo_act, fo_force, fo_contact;          // quad offs
  int32_t fo_tri;     // capsule_box rows: {p0, p1, p2, normal} of the row's triangle, 4 quads per row
  // state quad rows (each row = 32 env slots x float4, + 1 pad quad)
  // q_ct: one quad {pos, penetration} per contact row; q_cn: one quad {normal} per
  // collider pair; q_cd: dlambda of the rows, 4 per quad (pbd)
  int32_t q_act, q_js, q_ct, q_cn, q_cd, q_pre, q_ic, q_ia, q_ij;
  int32_t n_quads;
  int32_t n_zero;  // quads from q_ic to zero at the start of every step
  // per-env collider friction (bx_step_io.friction): float index in shared memory of the tile's 32
  // friction[e]^2 values (behind the mbarriers), set per launch; 0 = the pair tables' friction
  int32_t fr_off;
  // per-env joint sockets (bx_step_io.joint_offset): quad index in shared memory of the tile's [J][32] parent
  // offsets (behind the friction values, in front of the policy's buffers), set per launch; 0 = the joint tables'
  int32_t jo_off;
  // box_heightmap pairs: the height grids of the system, concatenated (device memory; NULL without one)
  const float* hm;
  BxSeg seg_in[BX_SEG_IN];
  BxSeg seg_out[BX_SEG_OUT];
  // scalars
  float dt, inv_dt, half_dt;
  float gravity[3];
  float vel_damp, ang_damp;
  float erp, collide_scale, vth;
  float inv_2h;  // 1 / (2*dt_sub), colliders.py:399
  // ---- env epilogue (optional; env_kind 0 = none): what Ant.step / Humanoid.step
  // compute after sys.step (envs/ant.py:222-281, envs/humanoid.py:246-340)
  int32_t env_kind;    // BX_ENV_*
  int32_t env_flags;   // bit 0 use_contact_forces, bit 1 terminate_when_unhealthy
  int32_t q_ep;        // persistent row: reference point BEFORE the step (torso pos / centre of mass)
  int32_t q_env;       // first env row (inside JS, after the angle_vel rows): layout EQ_* below
  float env_dt;        // config.dt
  float env_wf, env_wc, env_wcc, env_hr;  // forward / ctrl / contact weights, healthy_reward
  float env_zmin, env_zmax;
  float env_msum;      // humanoid: sum of the masses of bodies [0, nb-1) (float32 sum on the host)
  // fetch / grasp
  int32_t env_ref, env_target, env_object, env_hand, env_touch[4];
  int32_t q_eloc;      // egocentric rows of body b: q_eloc_base(b) + {0: pos_local, 1: vel_local}; pbd: the
                       // dead previous-pose rows of the body block (-1), else first row of 2 per body in JS
  int32_t q_ecf;       // contact flags, 4 bodies per quad
  int32_t q_ev;        // env vectors (EV_*)
  int32_t fo_envact;   // grasp action map in the float table: ceil(A/4) quads of min, then of range
  float env_radius, env_tdist, env_theight;
};

// env rows relative to q_env
enum {
  EQ_TORSO_POS = 0,  // {pos of body 0}
  EQ_TORSO_ROT = 1,  // {rot of body 0 (w,x,y,z)}
  EQ_S0 = 2,         // {reward, done, metric 0, metric 1}
  EQ_S1 = 3,         // {metric 2..5}
  EQ_S2 = 4,         // {metric 6..9}
  EQ_X = 5,          // ant: per body {clip(contact.vel)}, {clip(contact.ang)} (2 rows);
                     // humanoid: per body b < nb-1 {com_inr rows 0..2}, {com_vel}, {com_ang} (5 rows),
                     // then one row per actuator {act[idx] * strength}
};
// env vectors relative to q_ev (fetch / grasp)
enum {
  EV_0 = 0,  // fetch: torso_fwd                     grasp: {object_local_dir, |object_local|}
  EV_1 = 1,  // fetch: torso_up                      grasp: {target_local_dir, |target_local|}
  EV_2 = 2,  // fetch: {target_local_dir, |..|}      grasp: {hand_to_obj, heading_to_obj}
  EV_3 = 3,  //                                      grasp: {obj_to_target_dir, |obj_to_target|}
  EV_4 = 4,  //                                      grasp: {obj_heading_to_target, 0, 0, 0}
  EV_COUNT = 5,
};
#ifndef BX_ENV_NONE  // also in include/brax_b200.h
#define BX_ENV_NONE 0
#define BX_ENV_ANT 1
#define BX_ENV_HUMANOID 2
#define BX_ENV_FETCH 3
#define BX_ENV_GRASP 4
#endif

#define BX_QSTRIDE 33  // quads per row: 32 env slots + 1 pad (conflict-free transposing I/O)
#define BX_ENVS_PER_CTA 32
// body block: pbd {pos0 rot0 pos1 rot1 vel ang}, legacy_spring {pos rot vel ang}
#define BX_BODY_QUADS_PBD 6
#define BX_BODY_QUADS_SPRING 4

#endif  // BX_TABLES_H_
