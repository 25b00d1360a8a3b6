// bx_api.cu -- the C ABI of include/brax_b200.h over the fused step kernel.
//
// bx_system_create() flattens the static description of a system
// (what brax.System.__init__ builds, reference brax/physics/system.py:53-84)
// into the int/float tables of bx_tables.h and uploads them once.  bx_step()
// is one kernel launch on the caller's stream: no allocation, no host sync.
// There is no CPU fallback: without a CUDA device every call fails loudly.
#include <cuda_runtime.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "../../include/brax_b200.h"
#include "bx_params.h"
#include "bx_tables.h"

#define BX_HOST_STREAMS 3

struct bx_system {
  BxHeader h;
  int device;
  int32_t* d_tab_i;
  float4* d_tab_f;
  size_t smem_bytes;
  int warps;  // warps per CTA (W)
  bool warps_chosen;  // set by bx_system_configure / bx_system_autotune: the launch keeps it as is
  int jc, cc; // joint / contact class of the kernel variant (bx_params.h)
  bool fusable;  // the fused contact section is legal for this system
  int sm_count;
  int smem_optin;  // sharedMemPerBlockOptin of the device
  // scratch for bx_step_host / bx_rollout
  float* d_scratch;
  size_t scratch_floats;
  cudaStream_t streams[BX_HOST_STREAMS];
  bx_reset_desc* d_reset;  // device copy of desc.reset (bx_default_qp), or NULL
  int reset_dof;
  float* d_hm;  // height grids of the box_heightmap pairs, concatenated in pair order, or NULL
  BxResetMap reset_map;  // reset joint -> step joint (per-env joint sockets in bx_default_qp_offsets)
};

cudaError_t bx_reset_launch(const bx_reset_desc* d_tables, int nb, int64_t num_envs, const float* joint_angle,
                            const float* joint_velocity, const float* joint_offset, const BxResetMap& map,
                            float* pos, float* rot, float* vel, float* ang, cudaStream_t stream);

static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};
static long long* g_dbg = nullptr;  // BX_PHASE_TIMING builds: device buffer of 16 counters
#ifdef BX_PHASE_TIMING
extern "C" void bx_debug_set_buffer(long long* d) { g_dbg = d; }
#endif

static int fail(const std::string& m) {
  g_err = m;
  return 1;
}
#define CUDA_OK(x)                                                         \
  do {                                                                     \
    cudaError_t e_ = (x);                                                  \
    if (e_ != cudaSuccess)                                                 \
      return fail(std::string(#x) + ": " + cudaGetErrorString(e_));        \
  } while (0)

extern "C" const char* bx_last_error(void) { return g_err.c_str(); }
extern "C" int64_t bx_launch_count(void) { return g_launches.load(); }

static int validate(const bx_system_desc* d) {
  if (!d) return fail("desc is NULL");
  if (d->struct_size != sizeof(bx_system_desc))
    return fail("bx_system_desc size mismatch (ABI skew)");
  if (d->abi_version != BX_ABI_VERSION) return fail("bx_system_desc abi_version mismatch");
  if (d->dynamics_mode != BX_MODE_PBD && d->dynamics_mode != BX_MODE_LEGACY_SPRING)
    return fail("unknown dynamics_mode");
  if (d->num_bodies < 1 || d->num_bodies > BX_MAX_BODIES) return fail("num_bodies out of range");
  if (d->num_joints < 0 || d->num_joints > BX_MAX_JOINTS) return fail("num_joints out of range");
  if (d->num_actuators < 0 || d->num_actuators > BX_MAX_ACTUATORS)
    return fail("num_actuators out of range");
  if (d->num_forces < 0 || d->num_forces > BX_MAX_FORCES) return fail("num_forces out of range");
  if (d->num_pairs < 0 || d->num_pairs > BX_MAX_PAIRS) return fail("num_pairs out of range");
  if (d->num_collider_groups > BX_MAX_GROUPS) return fail("too many collider groups");
  if (d->act_size < 1) return fail("act_size must be >= 1");
  if (d->substeps < 1) return fail("substeps must be >= 1");
  if (!(d->dt_sub > 0.f)) return fail("dt_sub must be positive");
  if (!d->bodies) return fail("bodies is NULL");
  if (d->num_joints > 0 && !d->joints) return fail("joints is NULL");
  if (d->num_actuators > 0 && !d->actuators) return fail("actuators is NULL");
  if (d->num_forces > 0 && !d->forces) return fail("forces is NULL");
  if (d->num_pairs > 0 && !d->pairs) return fail("pairs is NULL");
  for (int i = 0; i < d->num_joints; i++) {
    const bx_joint& j = d->joints[i];
    if (j.kind < BX_JOINT_REVOLUTE || j.kind > BX_JOINT_SPRING_SPHERICAL)
      return fail("unsupported joint kind");
    bool spring = j.kind >= BX_JOINT_SPRING_REVOLUTE;
    if (spring != (d->dynamics_mode == BX_MODE_LEGACY_SPRING))
      return fail("joint kind does not match dynamics_mode");
    if (j.body_p < 0 || j.body_p >= d->num_bodies || j.body_c < 0 || j.body_c >= d->num_bodies)
      return fail("joint body index out of range");
  }
  for (int i = 0; i < d->num_actuators; i++) {
    const bx_actuator& a = d->actuators[i];
    if (a.kind != BX_ACT_TORQUE && a.kind != BX_ACT_ANGLE) return fail("unsupported actuator kind");
    if (a.joint < 0 || a.joint >= d->num_joints) return fail("actuator joint out of range");
    for (int k = 0; k < 3; k++)
      if (a.act_index[k] < -1 || a.act_index[k] >= d->act_size) return fail("actuator act_index out of range (-1 = masked)");
  }
  for (int i = 0; i < d->num_forces; i++) {
    const bx_force& f = d->forces[i];
    if (f.kind != BX_FORCE_THRUSTER && f.kind != BX_FORCE_TWISTER) return fail("unsupported force kind");
    if (f.body < 0 || f.body >= d->num_bodies) return fail("force body out of range");
    // (act_index is NOT range-checked: the reference reads it with jp.take(..., mode='clip'), forces.py:44,87,
    // and its own swimmer config indexes past the action vector; the kernel clamps the same way)
  }
  int k = 0;
  for (int i = 0; i < d->num_pairs; i++) {
    const bx_pair& p = d->pairs[i];
    if (p.kind < BX_PAIR_CAPSULE_PLANE || p.kind > BX_PAIR_BOX_HEIGHTMAP)
      return fail("unsupported collider pair kind (general meshes and box-box are out of scope)");
    if (p.body_a < 0 || p.body_a >= d->num_bodies || p.body_b < 0 || p.body_b >= d->num_bodies)
      return fail("pair body index out of range");
    if (p.group < 0 || p.group >= BX_MAX_GROUPS) return fail("pair group out of range");
    if (i > 0 && p.group < d->pairs[i - 1].group) return fail("pairs must be sorted by group");
    if (p.n_contacts < 1 || p.n_contacts > 12) return fail("pair n_contacts out of range");
    if (p.kind != BX_PAIR_CAPSULE_BOX && p.n_contacts > 8) return fail("pair n_contacts out of range");
    if (p.kind == BX_PAIR_CAPSULE_CAPSULE && p.n_contacts != 1) return fail("capsule_capsule has 1 contact");
    if (p.kind == BX_PAIR_CAPSULE_BOX && p.n_contacts != 12) return fail("capsule_box has 12 contacts (one per triangle)");
    if (p.kind != BX_PAIR_CAPSULE_CAPSULE && p.kind != BX_PAIR_CAPSULE_BOX && p.two_way)
      return fail("two-way plane colliders do not exist");
    if (p.kind == BX_PAIR_BOX_HEIGHTMAP && (!p.hm_height || p.hm_size < 2 || p.hm_size > 1024 || !(p.hm_cell_size > 0.f)))
      return fail("box_heightmap needs hm_height, 2 <= hm_size <= 1024 and hm_cell_size > 0");
    k += p.n_contacts;
  }
  if (k > BX_MAX_CONTACTS) return fail("too many contact rows");
  if (d->reset) {
    const bx_reset_desc& r = *d->reset;
    if (r.struct_size != sizeof(bx_reset_desc)) return fail("bx_reset_desc size mismatch (ABI skew)");
    if (r.num_dof < 0 || r.num_joints < 0 || r.num_joints > BX_MAX_JOINTS || r.num_probes < 0 ||
        r.num_probes > BX_MAX_PROBES || r.num_trees < 0 || r.num_trees > BX_MAX_TREES)
      return fail("bx_reset_desc counts out of range");
    for (int i = 0; i < r.num_joints; i++) {
      const bx_reset_joint& j = r.joints[i];
      if (j.body_p < 0 || j.body_p >= d->num_bodies || j.body_c < 0 || j.body_c >= d->num_bodies)
        return fail("bx_reset_desc joint body index out of range");
      for (int x = 0; x < 3; x++)
        if (j.take[x] < 0 || j.take[x] > r.num_dof) return fail("bx_reset_desc joint take index out of range");
    }
    for (int i = 0; i < r.num_probes; i++)
      if (r.probes[i].body < 0 || r.probes[i].body >= d->num_bodies || r.probes[i].tree < 0 ||
          r.probes[i].tree >= r.num_trees)
        return fail("bx_reset_desc probe index out of range");
  }
  if (d->env) {
    const bx_env_desc& e = *d->env;
    if (e.struct_size != sizeof(bx_env_desc)) return fail("bx_env_desc size mismatch (ABI skew)");
    if (e.kind < BX_ENV_ANT || e.kind > BX_ENV_GRASP)
      return fail("unsupported env kind (fused epilogues exist for ant, humanoid, fetch and grasp)");
    if (e.kind == BX_ENV_FETCH || e.kind == BX_ENV_GRASP) {
      const int idx[8] = {e.ref_body, e.target_body, e.kind == BX_ENV_GRASP ? e.object_body : 0,
                          e.kind == BX_ENV_GRASP ? e.hand_body : 0,
                          e.kind == BX_ENV_GRASP ? e.touch_bodies[0] : 0, e.kind == BX_ENV_GRASP ? e.touch_bodies[1] : 0,
                          e.kind == BX_ENV_GRASP ? e.touch_bodies[2] : 0, e.kind == BX_ENV_GRASP ? e.touch_bodies[3] : 0};
      for (int i = 0; i < 8; i++)
        if (idx[i] < 0 || idx[i] >= d->num_bodies) return fail("env body index out of range");
      if (e.kind == BX_ENV_GRASP && d->dynamics_mode != BX_MODE_PBD) return fail("grasp env epilogue: pbd only");
      if (e.kind == BX_ENV_GRASP && (d->act_size > BX_MAX_ACT || d->act_size < 3))
        return fail("grasp env epilogue: act_size out of range");
    }
    if (!(e.dt > 0.f)) return fail("env dt must be positive");
    if (e.kind == BX_ENV_HUMANOID && d->num_bodies < 2) return fail("humanoid epilogue needs >= 2 bodies");
    for (int i = 0; i < d->num_joints; i++)
      if (e.joint_free_dofs[i] < 0 || e.joint_free_dofs[i] > 3) return fail("joint_free_dofs out of range");
    for (int i = 0; i < d->num_actuators; i++)
      if (e.actuator_width[i] < 1 || e.actuator_width[i] > 3) return fail("actuator_width out of range");
  }
  return 0;
}

// The outputs of bx_step_io in segment order (pos_o is the first of BX_SEG_OUT pointers).
static_assert(offsetof(bx_step_io, env_rand) == (BX_SEG_IN + BX_SEG_OUT) * sizeof(void*), "bx_step_io vs segments");
static float* const* io_outs(const bx_step_io* io) { return &io->pos_o; }
static float** io_outs(bx_step_io* io) { return &io->pos_o; }

// Flattens the description into the tables of bx_tables.h.
static int joint_class(const bx_system_desc* d);

static void build_tables_g(const bx_system_desc* d, BxHeader* hh, std::vector<int32_t>* ti_out,
                           std::vector<float>* tf_out, int rows_per_item) {
  BxHeader& h = *hh;
  memset(&h, 0, sizeof(h));
  std::vector<int32_t>& ti = *ti_out;
  std::vector<float>& tf = *tf_out;  // 4 floats per quad
  ti.clear();
  tf.clear();
  const int nb = d->num_bodies, nj = d->num_joints, na = d->num_actuators, nf = d->num_forces;
  const bool spring = d->dynamics_mode == BX_MODE_LEGACY_SPRING;
  const int bqn = spring ? BX_BODY_QUADS_SPRING : BX_BODY_QUADS_PBD;
  h.mode = d->dynamics_mode;
  h.iters = spring ? d->substeps : d->substeps / 2;  // system.py:324 / :376
  h.nb = nb; h.nj = nj; h.na = na; h.nf = nf;
  h.act_size = d->act_size;
  auto quad = [&](float x, float y, float z, float w) {
    tf.push_back(x); tf.push_back(y); tf.push_back(z); tf.push_back(w);
  };
  auto nquads = [&]() { return static_cast<int>(tf.size() / 4); };
  auto align4 = [&]() { while (ti.size() & 3) ti.push_back(0); };

  // contact rows: pairs expanded, group-major / pair-major / contact-in-pair
  struct Row { int pair, c; };
  std::vector<Row> rows;
  for (int i = 0; i < d->num_pairs; i++)
    for (int c = 0; c < d->pairs[i].n_contacts; c++) rows.push_back({i, c});
  const int nk = static_cast<int>(rows.size());
  h.nk = nk;
  // row of the contact NORMAL of contact row k: one per collider pair (all rows of a plane pair share
  // the plane's normal; capsule_capsule has one row), one per ROW for capsule_box (a normal per triangle)
  std::vector<int> normal_row(nk, 0), tri_index(nk, -1);
  int n_normals = 0, n_tris = 0;
  for (int k = 0; k < nk; k++) {
    const int kind_k = d->pairs[rows[k].pair].kind;
    const bool per_row = kind_k == BX_PAIR_CAPSULE_BOX || kind_k == BX_PAIR_BOX_HEIGHTMAP;   // a normal per row
    const bool new_pair = k == 0 || rows[k].pair != rows[k - 1].pair;
    if (per_row || new_pair) n_normals++;
    normal_row[k] = n_normals - 1;
    if (kind_k == BX_PAIR_CAPSULE_BOX) tri_index[k] = n_tris++;
    // box_heightmap: the rows of a pair share one 4-quad block {cell_size, size, offset of the grid in h.hm}
    if (kind_k == BX_PAIR_BOX_HEIGHTMAP) {
      if (new_pair) tri_index[k] = n_tris++; else tri_index[k] = tri_index[k - 1];
    }
    // capsule_clipped_plane: the rows of a pair share one 8-quad block {normal, pos, x, y, halfsize, -, -, -}
    if (d->pairs[rows[k].pair].kind == BX_PAIR_CAPSULE_CLIPPED_PLANE) {
      if (new_pair) { tri_index[k] = n_tris; n_tris += 2; } else tri_index[k] = tri_index[k - 1];
    }
  }
  // contact items: up to rows_per_item consecutive rows of one pair share one
  // slot-list entry per side
  std::vector<int> item_k0, item_n, item_of(nk, 0);
  for (int k = 0; k < nk;) {
    int n = 1;
    while (n < rows_per_item && k + n < nk && rows[k + n].pair == rows[k].pair) n++;
    for (int r = 0; r < n; r++) item_of[k + r] = static_cast<int>(item_k0.size());
    item_k0.push_back(k);
    item_n.push_back(n);
    k += n;
  }
  const int nci = static_cast<int>(item_k0.size());
  h.nci = nci;

  // slots
  std::vector<int> cslot(nb, -1), aslot(nb, -1);
  int ncb = 0, nab = 0;
  for (int i = 0; i < d->num_pairs; i++) {
    const bx_pair& p = d->pairs[i];
    if (cslot[p.body_a] < 0) cslot[p.body_a] = ncb++;
    if (p.two_way && cslot[p.body_b] < 0) cslot[p.body_b] = ncb++;
  }
  for (int i = 0; i < na; i++) {
    const bx_joint& j = d->joints[d->actuators[i].joint];
    if (aslot[j.body_p] < 0) aslot[j.body_p] = nab++;
    if (aslot[j.body_c] < 0) aslot[j.body_c] = nab++;
  }
  h.ncb = ncb; h.nab = nab;

  // ---- float4 tables
  h.fo_body = nquads();
  for (int b = 0; b < nb; b++) {
    const bx_body& s = d->bodies[b];
    quad(1.f / s.mass, s.mass, 0.f, 0.f);
    quad(s.inv_inertia[0], s.inv_inertia[1], s.inv_inertia[2], 0.f);
    quad(s.pos_mask[0], s.pos_mask[1], s.pos_mask[2], 0.f);
    quad(s.rot_mask[0], s.rot_mask[1], s.rot_mask[2], 0.f);
  }
  h.fo_joint = nquads();
  // the revolute class's kernel variants read the 10-quad record (bx_step_kernel.cuh JQS)
  const bool rev_only = joint_class(d) == BX_JC_REVOLUTE;
  for (int i = 0; i < nj; i++) {
    const bx_joint& s = d->joints[i];
    quad(s.off_p[0], s.off_p[1], s.off_p[2], s.angular_damping);
    quad(s.off_c[0], s.off_c[1], s.off_c[2], s.scale_pos);
    const float wp[3] = {s.scale_ang, s.stiffness, s.spring_damping};
    const float wc[3] = {s.limit_strength, 0.f, 0.f};
    for (int r = 0; r < 3; r++) quad(s.axis_p[3 * r], s.axis_p[3 * r + 1], s.axis_p[3 * r + 2], wp[r]);
    for (int r = 0; r < 3; r++) quad(s.axis_c[3 * r], s.axis_c[3 * r + 1], s.axis_c[3 * r + 2], wc[r]);
    quad(s.limit[0], s.limit[1], s.limit[2], s.limit[3]);
    if (!rev_only) quad(s.limit[4], s.limit[5], 0.f, 0.f);
    for (int dd = 0; dd < (rev_only ? 1 : 3); dd++) {  // direction vectors of the limits (clip_dir, bx_step_kernel.cuh)
      float cs[4];
      for (int e = 0; e < 2; e++) {
        const double a = s.limit[2 * dd + e];
        const bool open_end = !(fabs(a) < 3.14159265358979323846 - 1e-6);  // a signed angle never passes +-pi
        cs[2 * e] = open_end ? -1.f : static_cast<float>(cos(a));
        cs[2 * e + 1] = open_end ? 0.f : static_cast<float>(sin(a));
      }
      quad(cs[0], cs[1], cs[2], cs[3]);
    }
  }
  h.fo_act = nquads();
  for (int i = 0; i < na; i++) quad(d->actuators[i].strength, 0.f, 0.f, 0.f);
  h.fo_force = nquads();
  for (int i = 0; i < nf; i++) quad(d->forces[i].strength, 0.f, 0.f, 0.f);
  h.fo_contact = nquads();
  for (int k = 0; k < nk; k++) {
    const bx_pair& p = d->pairs[rows[k].pair];
    quad(p.friction, p.elasticity, p.radius_a, p.radius_b);
    if (p.kind == BX_PAIR_CAPSULE_CAPSULE || p.kind == BX_PAIR_CAPSULE_BOX) {
      for (int q = 0; q < 4; q++) quad(p.pts[q][0], p.pts[q][1], p.pts[q][2], 0.f);
    } else {
      const float* pt = p.pts[rows[k].c];
      quad(pt[0], pt[1], pt[2], 0.f);
      for (int q = 0; q < 3; q++) quad(0.f, 0.f, 0.f, 0.f);
    }
  }
  // capsule_box rows: {p0, p1, p2, normal} of the row's triangle (body-b frame), 4 quads each
  h.fo_tri = nquads();
  for (int k = 0; k < nk; k++)
    if (tri_index[k] >= 0 && d->pairs[rows[k].pair].kind == BX_PAIR_BOX_HEIGHTMAP) {
      if (k > 0 && rows[k].pair == rows[k - 1].pair) continue;   // one block per pair
      const bx_pair& p = d->pairs[rows[k].pair];
      int off = 0;   // grids are concatenated in pair order (bx_system_create uploads them the same way)
      for (int q = 0; q < rows[k].pair; q++)
        if (d->pairs[q].kind == BX_PAIR_BOX_HEIGHTMAP) off += d->pairs[q].hm_size * d->pairs[q].hm_size;
      quad(p.hm_cell_size, static_cast<float>(p.hm_size), static_cast<float>(off), 0.f);
      for (int v = 0; v < 3; v++) quad(0.f, 0.f, 0.f, 0.f);
    } else if (tri_index[k] >= 0 && d->pairs[rows[k].pair].kind == BX_PAIR_CAPSULE_CLIPPED_PLANE) {
      if (k > 0 && rows[k].pair == rows[k - 1].pair) continue;   // one block per pair
      const bx_pair& p = d->pairs[rows[k].pair];
      for (int v = 0; v < 3; v++) quad(p.tri[0][v][0], p.tri[0][v][1], p.tri[0][v][2], 0.f);   // normal, pos, x
      quad(p.tri[1][0][0], p.tri[1][0][1], p.tri[1][0][2], 0.f);                              // y
      quad(p.tri[1][1][0], p.tri[1][1][1], 0.f, 0.f);                                          // halfsize_x, halfsize_y
      for (int v = 0; v < 3; v++) quad(0.f, 0.f, 0.f, 0.f);
    } else if (tri_index[k] >= 0) {
      const bx_pair& p = d->pairs[rows[k].pair];
      const int f = rows[k].c;
      for (int v = 0; v < 3; v++) quad(p.tri[f][v][0], p.tri[f][v][1], p.tri[f][v][2], 0.f);
      quad(p.tri_normal[f][0], p.tri_normal[f][1], p.tri_normal[f][2], 0.f);
    }
  if (d->env && d->env->kind == BX_ENV_GRASP) {  // action map: min quads, then range quads
    h.fo_envact = nquads();
    const int nq = (d->act_size + 3) / 4;
    for (int pass = 0; pass < 2; pass++)
      for (int q = 0; q < nq; q++) {
        float v[4];
        for (int x = 0; x < 4; x++) {
          const int i = 4 * q + x;
          v[x] = i < d->act_size ? (pass ? d->env->act_range[i] : d->env->act_min[i]) : 0.f;
        }
        quad(v[0], v[1], v[2], v[3]);
      }
  }
  h.n_fquads = nquads();

  // ---- per-body slot lists inside region JS (see bx_tables.h "Slots") ---------
  // Lists of one barrier interval are packed from quad 0; different intervals
  // reuse the region.  Order inside a list = the reference's summation order.
  std::vector<int> a_base(nb, 0), a_n(nb, 0), d_base(nb, 0), d_n(nb, 0), j_base(nb, 0), j_n(nb, 0);
  std::vector<int> act_slot_p(na), act_slot_c(na), dslot_p(nj), dslot_c(nj), jslot_p(nj), jslot_c(nj);
  int t_quads = 0;  // interval T: actuator lists, then (pbd) damping lists
  for (int b = 0; b < nb; b++) {
    a_base[b] = t_quads;
    // actuators.py:67-70: group by group, parents then children
    for (int i0 = 0; i0 < na;) {
      int i1 = i0;
      while (i1 < na && d->actuators[i1].group == d->actuators[i0].group) i1++;
      for (int i = i0; i < i1; i++)
        if (d->joints[d->actuators[i].joint].body_p == b) act_slot_p[i] = t_quads++;
      for (int i = i0; i < i1; i++)
        if (d->joints[d->actuators[i].joint].body_c == b) act_slot_c[i] = t_quads++;
      i0 = i1;
    }
    a_n[b] = t_quads - a_base[b];
  }
  if (!spring) {
    for (int b = 0; b < nb; b++) {
      d_base[b] = t_quads;
      // joints.py:93-98 / :103-128: group by group, parents then children
      for (int i0 = 0; i0 < nj;) {
        int i1 = i0;
        while (i1 < nj && d->joints[i1].group == d->joints[i0].group) i1++;
        for (int i = i0; i < i1; i++) if (d->joints[i].body_p == b) dslot_p[i] = t_quads++;
        for (int i = i0; i < i1; i++) if (d->joints[i].body_c == b) dslot_c[i] = t_quads++;
        i0 = i1;
      }
      d_n[b] = t_quads - d_base[b];
    }
  }
  // interval J (pbd: own interval from quad 0; spring: same interval as T, after it)
  int j_quads = spring ? t_quads : 0;
  for (int b = 0; b < nb; b++) {
    j_base[b] = j_quads;
    for (int i0 = 0; i0 < nj;) {
      int i1 = i0;
      while (i1 < nj && d->joints[i1].group == d->joints[i0].group) i1++;
      for (int i = i0; i < i1; i++) if (d->joints[i].body_p == b) { jslot_p[i] = j_quads; j_quads += 2; }
      for (int i = i0; i < i1; i++) if (d->joints[i].body_c == b) { jslot_c[i] = j_quads; j_quads += 2; }
      i0 = i1;
    }
    j_n[b] = (j_quads - j_base[b]) / 2;
  }
  // contact intervals: per body, per collider group, a rows then b rows
  // (colliders.py:222-228)
  std::vector<int> cslot_a(nci, 0), cslot_b(nci, 0);
  std::vector<std::vector<int>> cgroups(nb);  // (base, entries) pairs
  int c_quads = 0;
  auto pair_of_item = [&](int i) -> const bx_pair& { return d->pairs[rows[item_k0[i]].pair]; };
  for (int b = 0; b < nb; b++) {
    for (int i0 = 0; i0 < nci;) {
      int g = pair_of_item(i0).group, i1 = i0;
      while (i1 < nci && pair_of_item(i1).group == g) i1++;
      const int base = c_quads;
      for (int i = i0; i < i1; i++)
        if (pair_of_item(i).body_a == b) { cslot_a[i] = c_quads; c_quads += 2; }
      for (int i = i0; i < i1; i++)
        if (pair_of_item(i).two_way && pair_of_item(i).body_b == b) { cslot_b[i] = c_quads; c_quads += 2; }
      if (c_quads > base) { cgroups[b].push_back(base); cgroups[b].push_back((c_quads - base) / 2); }
      i0 = i1;
    }
  }
  int js = std::max(std::max(t_quads, j_quads), c_quads);
  const int cd_rows = spring ? 0 : (nk + 3) / 4;
  const int ct_rows = nk + n_normals + cd_rows;
  {
    // region JS doubles as the staging buffer of the bulk-copy I/O: it must hold
    // the largest output span of a tile, and JS..end must hold all input spans.
    const int row_floats = BX_QSTRIDE * 4;
    const int max_out = 32 * std::max(4 * nb, 3 * nk);
    js = std::max(js, (max_out + row_floats - 1) / row_floats);
    const int in_floats = 32 * (13 * nb + d->act_size);
    const int tail = ct_rows + (spring ? 0 : 2 * ncb) + 2 * ncb + nab + (spring ? 2 * nb : 0);
    js = std::max(js, (in_floats + row_floats - 1) / row_floats - tail);
    js = std::max(js, 2 * nj);  // angle_vel epilogue: {angles}, {velocities} per joint
    if (d->env) {  // env epilogue rows follow the angle_vel rows (bx_tables.h EQ_*)
      const int k = d->env->kind;
      int extra = k == BX_ENV_ANT ? 2 * nb : (k == BX_ENV_HUMANOID ? 5 * (nb - 1) + na : 0);
      if (k == BX_ENV_FETCH || k == BX_ENV_GRASP)  // flags + vectors (+ local rows where no dead body rows exist)
        extra = (nb + 3) / 4 + EV_COUNT + (spring ? 2 * nb : 0);
      js = std::max(js, 2 * nj + EQ_X + extra);
    }
  }

  // ---- state quad rows
  int r = nb * bqn;
  h.q_act = r; r += (d->act_size + 3) / 4;
  h.q_js = r; r += js;
  h.q_ct = r; r += nk;
  h.q_cn = r; r += n_normals;
  h.q_cd = r; r += cd_rows;
  h.q_pre = r; r += spring ? 0 : 2 * ncb;
  h.q_ic = r; r += 2 * ncb;   // q_ic, q_ia, q_ij contiguous: zeroed together
  h.q_ia = r; r += nab;
  h.q_ij = r; r += spring ? 2 * nb : 0;
  h.n_zero = r - h.q_ic;
  // survives the substeps: reference point before the step (grasp has none)
  h.q_ep = r; r += (d->env && d->env->kind != BX_ENV_GRASP) ? 1 : 0;
  h.q_env = h.q_js + 2 * nj;
  h.n_quads = r;

  // ---- int tables (every record 16-byte aligned)
  std::vector<int32_t> adj;
  std::vector<int32_t> body_rec(nb * BI_STRIDE, 0);
  for (int b = 0; b < nb; b++) {
    int32_t* rec = &body_rec[b * BI_STRIDE];
    rec[BI_CSLOT] = cslot[b];
    rec[BI_ASLOT] = aslot[b];
    rec[BI_A_BASE] = a_base[b]; rec[BI_A_N] = a_n[b];
    rec[BI_D_BASE] = d_base[b]; rec[BI_D_N] = d_n[b];
    rec[BI_J_BASE] = j_base[b]; rec[BI_J_N] = j_n[b];
    // forces: thrusters, then twisters (forces.py:131-136)
    rec[BI_F_OFF] = static_cast<int>(adj.size());
    for (int kind = 0; kind < 2; kind++)
      for (int i = 0; i < nf; i++)
        if (d->forces[i].kind == kind && d->forces[i].body == b) adj.push_back(i);
    rec[BI_F_N] = static_cast<int>(adj.size()) - rec[BI_F_OFF];
    rec[BI_CG_OFF] = static_cast<int>(adj.size());
    adj.insert(adj.end(), cgroups[b].begin(), cgroups[b].end());
    rec[BI_CG_N] = static_cast<int>(cgroups[b].size() / 2);
    rec[BI_ROW_OFF] = static_cast<int>(adj.size());
    for (int i = 0; i < nci; i++)
      if (pair_of_item(i).body_a == b) adj.push_back(i);
    rec[BI_ROW_N] = static_cast<int>(adj.size()) - rec[BI_ROW_OFF];
  }
  std::vector<int32_t> forder;
  for (int pass = 0; pass < 2; pass++)
    for (int b = 0; b < nb; b++)
      if ((body_rec[b * BI_STRIDE + BI_ROW_N] > 0) == (pass == 0)) forder.push_back(b);
  h.io_body = static_cast<int>(ti.size());
  ti.insert(ti.end(), body_rec.begin(), body_rec.end());
  align4();
  h.io_joint = static_cast<int>(ti.size());
  for (int i = 0; i < nj; i++) {
    ti.push_back(d->joints[i].kind);
    ti.push_back(d->joints[i].body_p);
    ti.push_back(d->joints[i].body_c);
    ti.push_back(0);
    ti.push_back(spring ? 0 : dslot_p[i]); ti.push_back(spring ? 0 : dslot_c[i]);
    ti.push_back(jslot_p[i]); ti.push_back(jslot_c[i]);
  }
  h.io_act = static_cast<int>(ti.size());
  for (int i = 0; i < na; i++) {
    const bx_actuator& a = d->actuators[i];
    ti.push_back(a.kind); ti.push_back(a.joint);
    for (int k = 0; k < 3; k++) ti.push_back(a.act_index[k]);
    ti.push_back(act_slot_p[i]); ti.push_back(act_slot_c[i]); ti.push_back(0);
  }
  h.io_force = static_cast<int>(ti.size());
  for (int i = 0; i < nf; i++) {
    const bx_force& f = d->forces[i];
    ti.push_back(f.kind); ti.push_back(f.body);
    for (int k = 0; k < 3; k++) ti.push_back(f.act_index[k]);
    for (int k = 0; k < 3; k++) ti.push_back(0);
  }
  h.io_contact = static_cast<int>(ti.size());
  for (int k = 0; k < nk; k++) {
    const bx_pair& p = d->pairs[rows[k].pair];
    ti.push_back(p.kind); ti.push_back(p.two_way); ti.push_back(p.body_a);
    ti.push_back(p.body_b); ti.push_back(cslot_a[item_of[k]]); ti.push_back(cslot_b[item_of[k]]);
    ti.push_back(normal_row[k]); ti.push_back(tri_index[k] < 0 ? 0 : 4 * tri_index[k]);
  }
  h.io_citem = static_cast<int>(ti.size());
  for (int i = 0; i < nci; i++) { ti.push_back(item_k0[i]); ti.push_back(item_n[i]); }
  align4();
  h.io_adj = static_cast<int>(ti.size());
  ti.insert(ti.end(), adj.begin(), adj.end());
  h.io_forder = static_cast<int>(ti.size());
  ti.insert(ti.end(), forder.begin(), forder.end());

  // ---- I/O segments: element r of an env -> shared float index of (quad row, comp)
  const int RF = BX_QSTRIDE * 4;  // floats per quad row
  auto seg = [&](int nper, float fill, const std::vector<int>& map) {
    BxSeg sgm;
    sgm.nper = nper;
    bool all_const = true;
    for (int v : map) all_const = all_const && v < 0;
    if (all_const) {  // constant segment: no map, the kernel does a plain fill
      sgm.map = -1;
      sgm.magic = 0;
      sgm.fill = fill;
      return sgm;
    }
    sgm.map = static_cast<int>(ti.size());
    sgm.magic = nper > 1 ? static_cast<uint32_t>((0x100000000ULL + nper - 1) / nper) : 0u;
    sgm.fill = fill;
    ti.insert(ti.end(), map.begin(), map.end());
    return sgm;
  };
  // per-body vector field with `w` components at quad (bqn*b + off); rot is stored (w,x,y,z)
  auto body_field = [&](int off, int w) {
    std::vector<int> m;
    for (int b = 0; b < nb; b++)
      for (int x = 0; x < w; x++) m.push_back((bqn * b + off) * RF + x);
    return m;
  };
  auto slot_field = [&](const std::vector<int>& slot, int base, int per, int off) {
    std::vector<int> m;
    for (int b = 0; b < nb; b++)
      for (int x = 0; x < 3; x++) m.push_back(slot[b] >= 0 ? (base + per * slot[b] + off) * RF + x : -1);
    return m;
  };
  const int vel_off = bqn - 2, ang_off = bqn - 1;
  h.seg_in[0] = seg(3 * nb, 0.f, body_field(0, 3));
  h.seg_in[1] = seg(4 * nb, 0.f, body_field(1, 4));
  h.seg_in[2] = seg(3 * nb, 0.f, body_field(vel_off, 3));
  h.seg_in[3] = seg(3 * nb, 0.f, body_field(ang_off, 3));
  {
    std::vector<int> m;
    for (int i = 0; i < d->act_size; i++) m.push_back((h.q_act + (i >> 2)) * RF + (i & 3));
    h.seg_in[4] = seg(d->act_size, 0.f, m);
  }
  std::vector<int> ident(nb), none(3 * nb, -1);
  for (int b = 0; b < nb; b++) ident[b] = b;
  h.seg_out[0] = h.seg_in[0];
  h.seg_out[1] = h.seg_in[1];
  h.seg_out[2] = h.seg_in[2];
  h.seg_out[3] = h.seg_in[3];
  h.seg_out[4] = seg(3 * nb, 0.f, slot_field(cslot, h.q_ic, 2, 0));   // contact.vel
  h.seg_out[5] = seg(3 * nb, 0.f, slot_field(cslot, h.q_ic, 2, 1));   // contact.ang
  h.seg_out[6] = seg(3 * nb, 0.f, spring ? slot_field(ident, h.q_ij, 2, 0) : none);  // joint.vel
  h.seg_out[7] = seg(3 * nb, 0.f, spring ? slot_field(ident, h.q_ij, 2, 1) : none);  // joint.ang
  h.seg_out[8] = seg(3 * nb, 0.f, none);                               // actuator.vel
  h.seg_out[9] = seg(3 * nb, 0.f, slot_field(aslot, h.q_ia, 1, 0));   // actuator.ang
  {
    std::vector<int> mp, mn, me;
    for (int k = 0; k < nk; k++) {
      for (int x = 0; x < 3; x++) {
        mp.push_back((h.q_ct + k) * RF + x);
        mn.push_back((h.q_cn + normal_row[k]) * RF + x);
      }
      me.push_back((h.q_ct + k) * RF + 3);
    }
    h.seg_out[10] = seg(3 * nk, 0.f, mp);
    h.seg_out[11] = seg(3 * nk, 0.f, mn);
    h.seg_out[12] = seg(nk, 0.f, me);
  }
  {  // angle_vel epilogue rows live in JS (dead after the last substep)
    std::vector<int> ma, mv;
    for (int i = 0; i < nj; i++)
      for (int x = 0; x < 3; x++) {
        ma.push_back((h.q_js + 2 * i) * RF + x);
        mv.push_back((h.q_js + 2 * i + 1) * RF + x);
      }
    h.seg_out[13] = seg(3 * nj, 0.f, ma);
    h.seg_out[14] = seg(3 * nj, 0.f, mv);
  }
  for (int i = BX_SEG_ENV; i < BX_SEG_OUT; i++) h.seg_out[i] = seg(0, 0.f, std::vector<int>());
  if (d->env) {
    // obs: what _get_obs concatenates (envs/ant.py:257-281, envs/humanoid.py:281-330),
    // every element an address of a row written by phase_angle_vel / phase_env
    const bx_env_desc& e = *d->env;
    const int E = h.q_env;
    std::vector<int> m;
    auto row3 = [&](int row, int x0 = 0, int n = 3) { for (int x = x0; x < x0 + n; x++) m.push_back(row * RF + x); };
    auto joints_part = [&](int which) {
      for (int i = 0; i < nj; i++) row3(h.q_js + 2 * i + which, 0, e.joint_free_dofs[i]);
    };
    const bool ego = e.kind == BX_ENV_FETCH || e.kind == BX_ENV_GRASP;
    int nmetrics = 10;
    if (ego) {
      // egocentric envs (fetch.py:98-122, grasp.py:144-186).  Rows: contact flags 4 per quad,
      // EV_* vectors, and per body {pos_local, vel_local}: in pbd these overwrite the body's
      // previous-pose rows (dead after the last substep: pos1/rot1 when the result sits in
      // pos0/rot0, which it does after the even number of half-steps of a full step), else JS rows
      nmetrics = 5;
      h.q_ecf = E + EQ_X;
      h.q_ev = h.q_ecf + (nb + 3) / 4;
      h.q_eloc = spring ? h.q_ev + EV_COUNT : -1;
      auto loc = [&](int b, int which) { return spring ? h.q_eloc + 2 * b + which : bqn * b + 2 + which; };
      auto flags = [&]() { for (int b = 0; b < nb; b++) m.push_back((h.q_ecf + (b >> 2)) * RF + (b & 3)); };
      auto locals = [&]() {
        for (int which = 0; which < 2; which++)
          for (int b = 0; b < nb; b++) row3(loc(b, which));
      };
      if (e.kind == BX_ENV_FETCH) {
        row3(h.q_ev + EV_0); row3(h.q_ev + EV_1);
        row3(h.q_ev + EV_2, 3, 1); row3(h.q_ev + EV_2);
        locals();
        flags();
      } else {
        row3(h.q_ev + EV_0, 3, 1); row3(h.q_ev + EV_0);
        row3(h.q_ev + EV_1, 3, 1); row3(h.q_ev + EV_1);
        locals();
        row3(h.q_ev + EV_2);                       // hand_to_obj
        row3(bqn * e.hand_body + vel_off);         // hand_vel
        row3(h.q_ev + EV_2, 3, 1);                 // heading_to_obj
        row3(h.q_ev + EV_3, 3, 1); row3(h.q_ev + EV_3);
        row3(h.q_ev + EV_4, 0, 1);
        flags();
      }
      h.env_ref = e.ref_body; h.env_target = e.target_body;
      h.env_object = e.object_body; h.env_hand = e.hand_body;
      for (int i = 0; i < 4; i++) h.env_touch[i] = e.touch_bodies[i];
      h.env_radius = e.target_radius;
      h.env_tdist = e.target_distance;
      h.env_theight = e.target_height;
    } else {
      if (e.exclude_current_positions) row3(E + EQ_TORSO_POS, 2, 1); else row3(E + EQ_TORSO_POS);
      row3(E + EQ_TORSO_ROT, 0, 4);
      joints_part(0);
      row3(bqn * 0 + vel_off);
      row3(bqn * 0 + ang_off);
      joints_part(1);
    }
    if (ego) {
    } else if (e.kind == BX_ENV_ANT) {
      if (e.use_contact_forces)
        for (int which = 0; which < 2; which++)
          for (int b = 0; b < nb; b++) row3(E + EQ_X + 2 * b + which);
    } else {
      nmetrics = 9;
      for (int b = 0; b < nb - 1; b++)
        for (int rr = 0; rr < 3; rr++) row3(E + EQ_X + 5 * b + rr);
      for (int b = 0; b < nb - 1; b++) row3(E + EQ_X + 5 * b + 3);
      for (int b = 0; b < nb - 1; b++) row3(E + EQ_X + 5 * b + 4);
      for (int i = 0; i < na; i++) row3(E + EQ_X + 5 * (nb - 1) + i, 0, e.actuator_width[i]);
    }
    h.seg_out[BX_SEG_ENV] = seg(static_cast<int>(m.size()), 0.f, m);
    h.seg_out[BX_SEG_ENV + 1] = seg(1, 0.f, std::vector<int>{(E + EQ_S0) * RF + 0});
    h.seg_out[BX_SEG_ENV + 2] = seg(1, 0.f, std::vector<int>{(E + EQ_S0) * RF + 1});
    std::vector<int> mm;
    for (int i = 0; i < nmetrics; i++) mm.push_back((E + EQ_S0) * RF + ((2 + i) >> 2) * RF + ((2 + i) & 3));
    h.seg_out[BX_SEG_ENV + 3] = seg(nmetrics, 0.f, mm);
    h.env_kind = e.kind;
    h.env_flags = (e.use_contact_forces ? 1 : 0) | (e.terminate_when_unhealthy ? 2 : 0) |
                  (ego ? 0 : 4);  // bit 2: the obs holds joint angles (phase_angle_vel runs)
    h.env_dt = e.dt;
    h.env_wf = e.forward_reward_weight;
    h.env_wc = e.ctrl_cost_weight;
    h.env_wcc = e.contact_cost_weight;
    h.env_hr = e.healthy_reward;
    h.env_zmin = e.healthy_z_min;
    h.env_zmax = e.healthy_z_max;
    float msum = 0.f;  // jp.sum(mass[:-1]) in float32
    for (int b = 0; b < nb - 1; b++) msum += d->bodies[b].mass;
    h.env_msum = msum;
  }
  align4();
  h.n_ints = static_cast<int>(ti.size());

  // ---- scalars
  h.dt = d->dt_sub;
  h.inv_dt = 1.f / d->dt_sub;
  h.half_dt = 0.5f * d->dt_sub;
  for (int k = 0; k < 3; k++) h.gravity[k] = d->gravity[k];
  h.vel_damp = d->vel_damp;
  h.ang_damp = d->ang_damp;
  h.erp = d->baumgarte_erp;
  h.collide_scale = d->collide_scale;
  h.vth = d->velocity_threshold;
  h.inv_2h = 1.f / (2.f * d->dt_sub);
}

static size_t tile_bytes(const BxHeader& h);

// Joint (and actuator) class of a system: which kernel variants run it (bx_params.h).
static int joint_class(const bx_system_desc* d) {
  bool rev = false, other = false, sph_only = d->num_joints > 0, angle = false, torque = false;
  for (int i = 0; i < d->num_joints; i++) {
    const int k = d->joints[i].kind;
    if (k == BX_JOINT_REVOLUTE || k == BX_JOINT_SPRING_REVOLUTE) rev = true; else other = true;
    if (k != BX_JOINT_SPHERICAL) sph_only = false;
  }
  for (int i = 0; i < d->num_actuators; i++)
    if (d->actuators[i].kind == BX_ACT_ANGLE) angle = true; else torque = true;
  if (d->num_forces > 0 || (angle && torque)) return BX_JC_ALL;
  if (!other) return angle ? BX_JC_ALL : BX_JC_REVOLUTE;
  if (sph_only && !rev) return angle ? BX_JC_SPHERICAL_ANGLE : BX_JC_SPHERICAL;
  return BX_JC_ALL;
}

// Rows per contact item: 1 (every row its own slot-list entry: most parallelism,
// the reference's exact summation order) unless pairing the rows is what lets two
// tiles share an SM (fetch: 8 box corners per collider).
static void build_tables(const bx_system_desc* d, BxHeader* h, std::vector<int32_t>* ti,
                         std::vector<float>* tf) {
  build_tables_g(d, h, ti, tf, 1);
  auto two_fit = [](const BxHeader& x) { return 2 * (tile_bytes(x) + 1024) <= 233472; };
  if (two_fit(*h)) return;
  BxHeader h2;
  std::vector<int32_t> ti2;
  std::vector<float> tf2;
  build_tables_g(d, &h2, &ti2, &tf2, 2);
  if (two_fit(h2)) { *h = h2; ti->swap(ti2); tf->swap(tf2); }
}

// The kernels live in one translation unit per launch bound (bx_kernels.cu); the
// launch picks the smallest bound that holds the configured CTA size.
static cudaError_t launch_variant(int mode, int jc, int cc, unsigned grid, int threads, size_t smem,
                                  cudaStream_t stream, const BxKernelParams& p, bool three) {
  // big batches of small systems: three tiles per SM (64 registers) when more than two waves
  // of tiles exist and three tiles fit in shared memory
  if ((cc & 7) == BX_CC_MESH)  // capsule-box systems: two builds exist (bx_kernels.cu BX_TU_MESH)
    return threads <= 320 ? bx_launch_320(mode, jc, cc, grid, threads, smem, stream, p)
                          : bx_launch_512(mode, jc, cc, grid, threads, smem, stream, p);
  if (three && threads <= 320 && 3 * (smem + 1024) <= 233472)
    return bx_launch_320x3(mode, jc, cc, grid, threads, smem, stream, p);
  if (threads <= 256) return bx_launch_256(mode, jc, cc, grid, threads, smem, stream, p);
  if (threads <= 320) return bx_launch_320(mode, jc, cc, grid, threads, smem, stream, p);
  // 12 warps with two tiles per SM (80 registers) only pays when two tiles fit in
  // shared memory; otherwise the one-tile-per-SM build with more registers is used
  if (threads <= 384 && 2 * (smem + 1024) <= 233472)
    return bx_launch_384(mode, jc, cc, grid, threads, smem, stream, p);
  return bx_launch_512(mode, jc, cc, grid, threads, smem, stream, p);
}
static int set_smem_attr(int bytes) {
  CUDA_OK(bx_set_smem_256(bytes));
  CUDA_OK(bx_set_smem_320(bytes));
  CUDA_OK(bx_set_smem_320x3(bytes));
  CUDA_OK(bx_set_smem_384(bytes));
  CUDA_OK(bx_set_smem_512(bytes));
  return 0;
}

static size_t tile_bytes(const BxHeader& h) {
  return (static_cast<size_t>(h.n_quads) * BX_QSTRIDE + h.n_fquads) * 16 +
         static_cast<size_t>(h.n_ints) * 4 + 16 /* mbarrier */ + BX_FRICTION_SMEM /* per-env friction */;
}

extern "C" int bx_desc_tile_bytes(const bx_system_desc* desc, int64_t* bytes, int32_t* rows) {
  if (validate(desc)) return 1;
  BxHeader h;
  std::vector<int32_t> ti;
  std::vector<float> tf;
  build_tables(desc, &h, &ti, &tf);
  if (bytes) *bytes = static_cast<int64_t>(tile_bytes(h));
  if (rows) {
    const int v[8] = {h.q_act, h.q_js - h.q_act, h.q_ct - h.q_js, h.q_pre - h.q_ct,
                      h.q_ic - h.q_pre, h.q_ia - h.q_ic, h.q_ij - h.q_ia, h.n_quads - h.q_ij};
    for (int i = 0; i < 8; i++) rows[i] = v[i];
  }
  return 0;
}

extern "C" int bx_system_create(const bx_system_desc* desc, int device, bx_system** out) {
  if (!out) return fail("out is NULL");
  *out = nullptr;
  if (validate(desc)) return 1;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail("no CUDA device: brax_b200 has no CPU fallback");
  if (device < 0 || device >= ndev) return fail("device index out of range");
  // the caller's current device is restored on every way out
  struct DeviceGuard {
    int prev = -1;
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  } guard;
  if (cudaGetDevice(&guard.prev) != cudaSuccess) guard.prev = -1;
  CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return fail("brax_b200 kernels are built for sm_100a (B200) only");

  bx_system* s = new bx_system();
  std::vector<int32_t> ti;
  std::vector<float> tf;
  build_tables(desc, &s->h, &ti, &tf);
  {
    s->jc = joint_class(desc);
    // fused contact section: pbd, one-way plane colliders only, <= 2 rows per body
    std::vector<int> rows_of(desc->num_bodies, 0);
    int max_rows = 0;
    for (int i = 0; i < desc->num_pairs; i++)
      max_rows = std::max(max_rows, rows_of[desc->pairs[i].body_a] += desc->pairs[i].n_contacts);
    int contact_bodies = 0;
    s->cc = BX_CC_PLANE;
    for (int b = 0; b < desc->num_bodies; b++) contact_bodies += rows_of[b] > 0;
    s->fusable = desc->dynamics_mode == BX_MODE_PBD && max_rows <= 2;
    for (int i = 0; i < desc->num_pairs; i++)
      if (desc->pairs[i].two_way || desc->pairs[i].kind == BX_PAIR_CAPSULE_CAPSULE ||
          desc->pairs[i].kind >= BX_PAIR_CAPSULE_BOX) {
        if (s->cc != BX_CC_MESH) s->cc = BX_CC_ALL;
        if (desc->pairs[i].kind >= BX_PAIR_CAPSULE_BOX) s->cc = BX_CC_MESH;   // capsule_box, capsule_clipped_plane, box_heightmap
        s->fusable = false;
      }
    // Measured (profiles/r01_sweeps.txt): the fused section wins when the chains
    // leave warps free (ant: 4 of 10 bodies collide, +2.5%) and loses when nearly
    // every body carries one (humanoid, -1%); bx_system_autotune times both.
    if (s->cc != BX_CC_ALL && s->cc != BX_CC_MESH)
      s->cc = (s->fusable && 2 * contact_bodies <= desc->num_bodies) ? BX_CC_PLANE_FUSED : BX_CC_PLANE;
  }
  s->device = device;
  s->warps_chosen = false;
  s->sm_count = prop.multiProcessorCount;
  s->smem_optin = static_cast<int>(prop.sharedMemPerBlockOptin);
  s->d_scratch = nullptr;
  s->scratch_floats = 0;
  for (int i = 0; i < BX_HOST_STREAMS; i++) s->streams[i] = nullptr;
  s->smem_bytes = tile_bytes(s->h);
  if (s->smem_bytes > static_cast<size_t>(prop.sharedMemPerBlockOptin)) {
    delete s;
    return fail("system state does not fit in shared memory (too many bodies/joints/contacts)");
  }
  // default warps per tile: minimise the critical path of one half-substep,
  // sum over phases of ceil(items / W) * (measured cycles of one item, in k-cycles,
  // tools/phase_timing.py), times a penalty for the register budget that W implies
  // (10 warps: 96 regs; 12: 80 regs; 16: one tile per SM); profiles/r01_sweeps.txt.
  {
    bool spherical = false;
    for (int i = 0; i < desc->num_joints; i++)
      spherical = spherical || (desc->joints[i].kind != BX_JOINT_REVOLUTE &&
                                desc->joints[i].kind != BX_JOINT_SPRING_REVOLUTE);
    const double cj = spherical ? 6.7 : 2.3, ca = spherical ? 4.0 : 1.7, cb = 3.56, ck = 1.07;
    const int cand[5] = {4, 8, 10, 12, 16};
    // 12 warps: two tiles per SM at 80 registers when shared memory allows (spherical
    // joint code spills there: humanoid -4%), else the one-tile-per-SM build
    const bool two_fit = 2 * (s->smem_bytes + 1024) <= 233472;
    const double pen[5] = {1.0, 1.0, 1.03, two_fit ? (spherical ? 1.3 : 1.06) : 2.0, 1.7};
    double best = 1e30;
    s->warps = 8;
    auto rounds = [](int n, int w) { return (n + w - 1) / w; };
    for (int i = 0; i < 5; i++) {
      const int w = cand[i];
      const double cost = pen[i] * (rounds(s->h.nj, w) * cj + rounds(s->h.na, w) * ca +
                                    rounds(s->h.nb, w) * cb +
                                    rounds(s->h.nci, w) * ck * (s->h.nci ? s->h.nk / s->h.nci : 1));
      if (cost < best - 1e-9) { best = cost; s->warps = w; }
    }
  }
  s->d_tab_i = nullptr;
  s->d_tab_f = nullptr;
  s->d_reset = nullptr;
  s->d_hm = nullptr;
  s->reset_dof = 0;
  auto drop = [&](const char* what, cudaError_t err) {   // frees whatever exists of a half-built system
    const std::string m = std::string(what) + (err != cudaSuccess ? std::string(": ") + cudaGetErrorString(err) : "");
    cudaFree(s->d_tab_i); cudaFree(s->d_tab_f); cudaFree(s->d_reset); cudaFree(s->d_hm);
    delete s;
    return fail(m);
  };
  cudaError_t e1 = cudaMalloc(&s->d_tab_i, std::max<size_t>(ti.size(), 1) * 4);
  cudaError_t e2 = cudaMalloc(&s->d_tab_f, std::max<size_t>(tf.size(), 1) * 4);
  if (e1 != cudaSuccess || e2 != cudaSuccess) return drop("cudaMalloc of tables failed", e1 != cudaSuccess ? e1 : e2);
  if ((e1 = cudaMemcpy(s->d_tab_i, ti.data(), ti.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e1 = cudaMemcpy(s->d_tab_f, tf.data(), tf.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess)
    return drop("upload of tables failed", e1);
  if (set_smem_attr(static_cast<int>(prop.sharedMemPerBlockOptin))) {
    const std::string keep = g_err;
    return drop(keep.c_str(), cudaSuccess);
  }
  {
    std::vector<float> hm;   // height grids of the box_heightmap pairs, in pair order
    for (int i = 0; i < desc->num_pairs; i++)
      if (desc->pairs[i].kind == BX_PAIR_BOX_HEIGHTMAP)
        hm.insert(hm.end(), desc->pairs[i].hm_height, desc->pairs[i].hm_height + desc->pairs[i].hm_size * desc->pairs[i].hm_size);
    if (!hm.empty()) {
      if ((e1 = cudaMalloc(&s->d_hm, hm.size() * 4)) != cudaSuccess) return drop("cudaMalloc of the height grids failed", e1);
      if ((e1 = cudaMemcpy(s->d_hm, hm.data(), hm.size() * 4, cudaMemcpyHostToDevice)) != cudaSuccess)
        return drop("upload of the height grids failed", e1);
      s->h.hm = s->d_hm;
    }
  }
  if (desc->reset) {
    if ((e1 = cudaMalloc(&s->d_reset, sizeof(bx_reset_desc))) != cudaSuccess) return drop("cudaMalloc of reset tables failed", e1);
    if ((e1 = cudaMemcpy(s->d_reset, desc->reset, sizeof(bx_reset_desc), cudaMemcpyHostToDevice)) != cudaSuccess)
      return drop("upload of reset tables failed", e1);
    s->reset_dof = desc->reset->num_dof;
    // reset joints are listed in depth order, the step's joints in the reference's joint-group order: match by bodies
    s->reset_map.num_joints = desc->num_joints;
    for (int j = 0; j < desc->reset->num_joints && j < 64; j++) {
      s->reset_map.step_joint[j] = -1;
      for (int k = 0; k < desc->num_joints; k++)
        if (desc->joints[k].body_p == desc->reset->joints[j].body_p && desc->joints[k].body_c == desc->reset->joints[j].body_c)
          s->reset_map.step_joint[j] = k;
    }
  }
  *out = s;
  return 0;
}

extern "C" int bx_system_num_joint_dof(const bx_system* s) { return s && s->d_reset ? s->reset_dof : -1; }

extern "C" int bx_default_qp(const bx_system* s, int64_t B, const float* joint_angle, const float* joint_velocity,
                             float* pos, float* rot, float* vel, float* ang, void* stream) {
  return bx_default_qp_offsets(s, B, joint_angle, joint_velocity, nullptr, pos, rot, vel, ang, stream);
}

extern "C" int bx_default_qp_offsets(const bx_system* s, int64_t B, const float* joint_angle, const float* joint_velocity,
                                     const float* joint_offset, float* pos, float* rot, float* vel, float* ang,
                                     void* stream) {
  if (!s) return fail("system is NULL");
  if (joint_offset)
    for (int j = 0; s->d_reset && j < s->reset_map.num_joints && j < 64; j++)
      if (s->reset_map.step_joint[j] < 0) return fail("bx_default_qp_offsets: a reset joint has no counterpart in desc.joints");
  if (!s->d_reset) return fail("bx_default_qp needs a system created with desc.reset");
  if (B <= 0) return fail("num_envs must be positive");
  if (!pos || !rot || !vel || !ang) return fail("QP pointers must not be NULL");
  if (s->reset_dof > 0 && (!joint_angle || !joint_velocity)) return fail("joint_angle / joint_velocity must not be NULL");
  CUDA_OK(cudaSetDevice(s->device));
  CUDA_OK(bx_reset_launch(s->d_reset, s->h.nb, B, joint_angle, joint_velocity, joint_offset, s->reset_map, pos, rot, vel,
                          ang, static_cast<cudaStream_t>(stream)));
  g_launches.fetch_add(1);
  return 0;
}

extern "C" void bx_system_destroy(bx_system* s) {
  if (!s) return;
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  struct Restore { int d; ~Restore() { if (d >= 0) cudaSetDevice(d); } } restore{prev};
  cudaSetDevice(s->device);
  cudaFree(s->d_tab_i);
  cudaFree(s->d_tab_f);
  cudaFree(s->d_scratch);
  cudaFree(s->d_reset);
  cudaFree(s->d_hm);
  for (int i = 0; i < BX_HOST_STREAMS; i++)
    if (s->streams[i]) cudaStreamDestroy(s->streams[i]);
  delete s;
}

extern "C" int bx_system_num_contacts(const bx_system* s) { return s ? s->h.nk : -1; }
extern "C" int bx_system_num_bodies(const bx_system* s) { return s ? s->h.nb : -1; }
extern "C" int bx_system_act_size(const bx_system* s) { return s ? s->h.act_size : -1; }
extern "C" int bx_system_obs_size(const bx_system* s) { return s ? s->h.seg_out[BX_SEG_ENV].nper : -1; }
extern "C" int bx_system_num_metrics(const bx_system* s) { return s ? s->h.seg_out[BX_SEG_ENV + 3].nper : -1; }

extern "C" int bx_system_configure(bx_system* s, int32_t threads_per_block, int32_t contact_section) {
  if (!s) return fail("system is NULL");
  if (threads_per_block != 0 &&
      (threads_per_block < 32 || threads_per_block > 512 || (threads_per_block & 31)))
    return fail("threads_per_block must be 0 or a multiple of 32 in [32, 512]");
  if (contact_section < 0 || contact_section > 2) return fail("contact_section must be 0, 1 or 2");
  if (contact_section == 2 && !s->fusable)
    return fail("fused contact section needs pbd, one-way plane colliders, <= 2 rows per body");
  if (threads_per_block) { s->warps = threads_per_block / 32; s->warps_chosen = true; }
  if (contact_section && s->cc != BX_CC_ALL && s->cc != BX_CC_MESH)
    s->cc = contact_section == 2 ? BX_CC_PLANE_FUSED : BX_CC_PLANE;
  return 0;
}

static int launch(const bx_system* s, int64_t num_envs, int32_t num_steps,
                  const bx_step_io* io, cudaStream_t stream, uint32_t stagger_ns = 0) {
  if (!s || !io) return fail("NULL argument");
  if (num_envs <= 0) return fail("num_envs must be positive");
  if (num_steps < 1) return fail("num_steps must be >= 1");
  if (!io->pos || !io->rot || !io->vel || !io->ang || !io->pos_o || !io->rot_o || !io->vel_o || !io->ang_o)
    return fail("QP / act pointers must not be NULL");
  if (!io->act && !io->policy) return fail("QP / act pointers must not be NULL");
  BxKernelParams p;
  memset(&p.pol, 0, sizeof(p.pol));
  p.obs0 = nullptr;
  p.act_out = nullptr;
  p.friction = io->friction;
  p.joint_offset = io->joint_offset;
  // per-env joint sockets: [J][32] quads behind the friction values (16-byte aligned: + 16 bytes of slack)
  const size_t jo_bytes = io->joint_offset ? 16 + static_cast<size_t>(s->h.nj) * BX_ENVS_PER_CTA * 16 : 0;
  const size_t base_bytes = s->smem_bytes + jo_bytes;   // the tile without the policy's buffers
  if (base_bytes > static_cast<size_t>(s->smem_optin)) return fail("the per-env joint offsets do not fit in shared memory beside the tile");
  size_t policy_smem = 0;
  bool policy_one_tile = false;   // the policy's buffers leave room for one tile per SM only
  if (io->policy) {
    // in-kernel policy (include/brax_b200.h bx_policy): validated here, passed by value
    const bx_policy& q = *io->policy;
    if (q.struct_size != sizeof(bx_policy)) return fail("bx_policy size mismatch (ABI skew)");
    if (!s->h.env_kind) return fail("the in-kernel policy needs a system created with desc.env (it reads the obs)");
    if (io->observe) return fail("policy and observe exclude each other");
    if (io->action_repeat > 1) return fail("policy with action_repeat > 1 is not supported");
    if (q.num_layers < 1 || q.num_layers > BX_POLICY_MAX_LAYERS) return fail("bx_policy.num_layers out of range");
    if (q.sizes[0] != s->h.seg_out[BX_SEG_ENV].nper) return fail("bx_policy.sizes[0] must be the obs size");
    int hmax = 0;
    for (int l = 1; l <= q.num_layers; l++) {
      if (q.sizes[l] < 1 || q.sizes[l] > 256) return fail("bx_policy layer width out of range (1..256)");
      hmax = std::max(hmax, q.sizes[l]);
    }
    const int outn = q.sizes[q.num_layers];
    if (outn != s->h.act_size && outn != 2 * s->h.act_size) return fail("bx_policy output width must be A or 2A");
    if (q.noise && outn != 2 * s->h.act_size) return fail("bx_policy.noise needs a 2A-wide output (loc, scale)");
    if (q.activation < BX_POLICY_RELU || q.activation > BX_POLICY_TANH) return fail("bx_policy.activation unknown");
    if (!q.params || !io->obs0) return fail("bx_policy.params and obs0 must not be NULL");
    p.pol.num_layers = q.num_layers;
    for (int l = 0; l <= q.num_layers; l++) p.pol.sizes[l] = q.sizes[l];
    p.pol.activation = q.activation;
    p.pol.hmax = hmax;
    p.pol.min_std = q.min_std;
    p.pol.params = q.params; p.pol.obs_mean = q.obs_mean; p.pol.obs_std = q.obs_std; p.pol.noise = q.noise;
    p.obs0 = io->obs0;
    p.act_out = io->act_out;
    // the normalised obs is staged whole when that keeps the tile's occupancy, else in chunks of 128 / 64 /
    // 32 rows (two more barriers per chunk): the largest that still lets two tiles share an SM, or fits at all
    const bool two_now = 2 * (base_bytes + 1024) <= 233472;
    const int cand[4] = {q.sizes[0], 128, 64, 32};
    int rows = 0;
    for (int pass = two_now ? 0 : 1; pass < 2 && !rows; pass++)
      for (int i = 0; i < 4 && !rows; i++) {
        if (cand[i] > q.sizes[0]) continue;
        const size_t total = base_bytes + BX_POLICY_SMEM(cand[i], hmax);
        if (pass == 0 ? 2 * (total + 1024) <= 233472 : total <= static_cast<size_t>(s->smem_optin)) rows = cand[i];
      }
    if (!rows) return fail("the policy's activation buffers do not fit in shared memory beside the tile");
    p.pol.obs_rows = rows;
    policy_smem = BX_POLICY_SMEM(rows, hmax);
    policy_one_tile = 2 * (base_bytes + policy_smem + 1024) > 233472;
  }
  p.h = s->h;
  // (float index behind the two mbarriers; 0 = off)
  p.h.fr_off = io->friction ? (s->h.n_quads * BX_QSTRIDE + s->h.n_fquads) * 4 + s->h.n_ints + 4 : 0;
  p.h.jo_off = io->joint_offset && s->h.nj > 0
                   ? s->h.n_quads * BX_QSTRIDE + s->h.n_fquads + ((s->h.n_ints * 4 + 16 + BX_FRICTION_SMEM + 15) >> 4)
                   : 0;
  p.tab_i = s->d_tab_i;
  p.tab_f = s->d_tab_f;
  p.in[0] = io->pos; p.in[1] = io->rot; p.in[2] = io->vel; p.in[3] = io->ang; p.in[4] = io->act;
  for (int i = 0; i < BX_SEG_OUT; i++) p.out[i] = io_outs(io)[i];
  if (!s->h.env_kind && (io->obs || io->reward || io->done || io->metrics))
    return fail("obs/reward/done/metrics need a system created with desc.env");
  if (io->env_rand && s->h.env_kind != BX_ENV_FETCH && s->h.env_kind != BX_ENV_GRASP)
    return fail("env_rand is the target teleport of the fetch / grasp epilogues");
  p.env_rand = io->env_rand;
  p.first[0] = io->first_pos; p.first[1] = io->first_rot; p.first[2] = io->first_vel; p.first[3] = io->first_ang;
  p.first_obs = io->first_obs;
  p.done_in = io->done_in; p.steps_in = io->steps_in;
  p.steps_out = io->steps_out; p.truncation = io->truncation;
  p.episode_length = io->episode_length;
  p.observe = io->observe;
  if (io->observe) {
    if (!s->h.env_kind) return fail("observe needs a system created with desc.env");
    if (num_steps != 1) return fail("observe is a single launch (bx_step)");
    if (!io->obs) return fail("observe needs the obs output");
    if (io->first_pos) return fail("observe and auto-reset exclude each other");
  }
  if (io->first_pos) {
    if (!s->h.env_kind) return fail("auto-reset needs a system created with desc.env");
    if (!io->first_rot || !io->first_vel || !io->first_ang || !io->done_in)
      return fail("auto-reset needs first_pos/rot/vel/ang and done_in");
    if (!io->done) return fail("auto-reset needs the done output");
    if (io->obs && !io->first_obs) return fail("auto-reset with an obs output needs first_obs");
  }
  const int repeat = io->action_repeat > 1 ? io->action_repeat : 1;
  if (repeat > 1) {
    if (!io->first_pos) return fail("action_repeat belongs to EpisodeWrapper: it needs the auto-reset inputs");
    if (s->h.env_kind == BX_ENV_GRASP) return fail("action_repeat is not available for grasp (its action map runs per launch step)");
    if (static_cast<int64_t>(num_steps) * repeat > 0x7fffffffLL) return fail("num_steps * action_repeat overflows");
  }
  p.action_repeat = repeat;
  p.num_envs = num_envs;
  p.num_steps = num_steps * repeat;
  p.stagger_ns = stagger_ns;
  p.dbg = g_dbg;
  const int64_t tiles = (num_envs + BX_ENVS_PER_CTA - 1) / BX_ENVS_PER_CTA;
  if (tiles > 0x7fffffffLL) return fail("too many environments for one launch");
  // systems with an env epilogue run the builds that contain it; all others the plain ones
  // BX_THREE_TILES: 0 never, 1 always (where it fits), default: batches of more than 4 tiles per SM
  static const int three_mode = getenv("BX_THREE_TILES") ? atoi(getenv("BX_THREE_TILES")) : 2;
  const bool three = three_mode == 1 || (three_mode == 2 && tiles > 4 * static_cast<int64_t>(s->sm_count));
  // BX_PDL=0 switches programmatic dependent launch off (A/B, debugging)
  // Not for grids of a quarter of a wave to one wave (one tile per SM): the early CTAs of launch
  // i+1 park in the free slots, and launch i+2 then finds an uneven slot map and doubles tiles
  // up on some SMs (measured, ant: 2048 envs 53.9 vs 49.2 us per step, 4096 envs 62.6 vs 51.9;
  // 1024 envs 46.0 vs 48.1 and everything above one wave gains).
  static const int pdl_mode = getenv("BX_PDL") ? atoi(getenv("BX_PDL")) : 2;
  const bool one_wave = tiles <= s->sm_count && 4 * tiles > s->sm_count;
  const int pdl = (pdl_mode == 1 || (pdl_mode == 2 && !one_wave)) ? BX_CC_PDL : 0;
  // the launch goes to the system's device whatever the caller's current device is (and leaves the
  // caller's current device as it was); `stream` must belong to the system's device
  int cur_dev = s->device;
  if (cudaGetDevice(&cur_dev) != cudaSuccess) cur_dev = s->device;
  if (cur_dev != s->device) CUDA_OK(cudaSetDevice(s->device));
  // At most one tile per SM (batches up to 32 x SM count: 4736 envs): nothing shares the SM, so the tile
  // takes 16 warps with the one-tile register budget -- every phase's items in one round (measured at
  // 4096 envs: grasp 38.9 -> 34.9 us, humanoid 55.3 -> 49.8, fetch 34.2 -> 30.8, ant 44.0 -> 43.0).
  // An explicit bx_system_configure / autotune choice is kept.
  // The same holds when the in-kernel policy's buffers leave room for one tile per SM only (humanoid).
  const int warps = (!s->warps_chosen && (tiles <= s->sm_count || policy_one_tile) && s->warps < 16) ? 16 : s->warps;
  // per-env joint sockets run the all-contact-features build of the system's joint class (BXF_JOINT_OFFSET, bx_kernels.cu)
  const int jc = s->jc, cc = p.h.jo_off ? (BX_CC_MESH | BX_CC_JO) : s->cc;
  cudaError_t le = launch_variant(s->h.mode, jc, cc | (s->h.env_kind ? BX_CC_ENV : 0) | pdl,
                                  static_cast<unsigned>(tiles), 32 * warps,
                                  base_bytes + policy_smem, stream, p, three);
  if (cur_dev != s->device) cudaSetDevice(cur_dev);
  g_launches.fetch_add(1);
  CUDA_OK(le);
  return 0;
}

extern "C" int bx_step(const bx_system* s, int64_t num_envs, const bx_step_io* io, void* stream) {
  return launch(s, num_envs, 1, io, static_cast<cudaStream_t>(stream));
}

extern "C" int bx_rollout(bx_system* s, int64_t num_envs, int32_t num_steps, const bx_step_io* io,
                          void* stream) {
  return launch(s, num_envs, num_steps, io, static_cast<cudaStream_t>(stream));
}

// ---- opaque device-resident state (include/brax_b200.h "bx_state") ------------------------------
struct bx_state {
  const bx_system* sys;
  int64_t num_envs;
  float* buf;  // one allocation: pos[B,N,3] rot[B,N,4] vel[B,N,3] ang[B,N,3], each on a 256-byte boundary
  float *pos, *rot, *vel, *ang;
};

extern "C" int bx_state_create(const bx_system* s, int64_t B, bx_state** out) {
  if (!out) return fail("out is NULL");
  *out = nullptr;
  if (!s) return fail("system is NULL");
  if (B <= 0) return fail("num_envs must be positive");
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  struct Restore { int d; ~Restore() { if (d >= 0) cudaSetDevice(d); } } restore{prev};
  CUDA_OK(cudaSetDevice(s->device));
  const size_t n = static_cast<size_t>(s->h.nb);
  auto pad = [](size_t floats) { return (floats + 63) / 64 * 64; };
  const size_t f3 = pad(3 * n * B), f4 = pad(4 * n * B);
  bx_state* st = new bx_state();
  st->sys = s;
  st->num_envs = B;
  st->buf = nullptr;
  cudaError_t e = cudaMalloc(&st->buf, (3 * f3 + f4) * sizeof(float));
  if (e != cudaSuccess) { delete st; return fail(std::string("cudaMalloc of the state failed: ") + cudaGetErrorString(e)); }
  st->pos = st->buf; st->rot = st->pos + f3; st->vel = st->rot + f4; st->ang = st->vel + f3;
  *out = st;
  return 0;
}
extern "C" void bx_state_destroy(bx_state* st) {
  if (!st) return;
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  cudaSetDevice(st->sys->device);
  cudaFree(st->buf);
  if (prev >= 0) cudaSetDevice(prev);
  delete st;
}
extern "C" int64_t bx_state_num_envs(const bx_state* st) { return st ? st->num_envs : -1; }
static int state_copy(const bx_state* st, float* const dst[4], const float* const src[4], cudaStream_t stream) {
  const size_t n = static_cast<size_t>(st->sys->h.nb) * st->num_envs;
  const size_t w[4] = {3, 4, 3, 3};
  for (int i = 0; i < 4; i++) {
    if (!dst[i] || !src[i]) return fail("QP pointers must not be NULL");
    CUDA_OK(cudaMemcpyAsync(dst[i], src[i], w[i] * n * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  }
  return 0;
}
extern "C" int bx_state_set(bx_state* st, const float* pos, const float* rot, const float* vel, const float* ang,
                            void* stream) {
  if (!st) return fail("state is NULL");
  float* const dst[4] = {st->pos, st->rot, st->vel, st->ang};
  const float* const src[4] = {pos, rot, vel, ang};
  return state_copy(st, dst, src, static_cast<cudaStream_t>(stream));
}
extern "C" int bx_state_get(const bx_state* st, float* pos, float* rot, float* vel, float* ang, void* stream) {
  if (!st) return fail("state is NULL");
  float* const dst[4] = {pos, rot, vel, ang};
  const float* const src[4] = {st->pos, st->rot, st->vel, st->ang};
  return state_copy(st, dst, src, static_cast<cudaStream_t>(stream));
}
extern "C" int bx_state_rollout(bx_state* st, int32_t num_steps, const bx_step_io* io, void* stream) {
  if (!st || !io) return fail("NULL argument");
  if (io->pos || io->rot || io->vel || io->ang || io->pos_o || io->rot_o || io->vel_o || io->ang_o)
    return fail("bx_state_rollout: the QP pointers of bx_step_io must be NULL (the state handle owns the QP)");
  bx_step_io d = *io;
  d.pos = st->pos; d.rot = st->rot; d.vel = st->vel; d.ang = st->ang;
  d.pos_o = st->pos; d.rot_o = st->rot; d.vel_o = st->vel; d.ang_o = st->ang;   // in place
  return launch(st->sys, st->num_envs, num_steps, &d, static_cast<cudaStream_t>(stream));
}

// Times the candidate CTA sizes on the caller's own buffers and keeps the fastest
// (outputs are overwritten by every trial: call it before a real step).
extern "C" int bx_system_autotune(bx_system* s, int64_t num_envs, const bx_step_io* io, void* stream) {
  if (!s || !io) return fail("NULL argument");
  CUDA_OK(cudaSetDevice(s->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  struct Events {
    cudaEvent_t a = nullptr, b = nullptr;
    ~Events() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
  } ev;
  CUDA_OK(cudaEventCreate(&ev.a));
  CUDA_OK(cudaEventCreate(&ev.b));
  const cudaEvent_t e0 = ev.a, e1 = ev.b;
  const int cand[4] = {8, 10, 12, 16};
  const int keep = s->warps, keep_cc = s->cc;
  s->warps_chosen = true;  // the trial launches run the candidate W as is
  float best = 1e30f;
  int best_w = keep, best_cc = keep_cc;
  // two interleaved passes over all candidates, 8 timed launches each, best pass
  // per candidate: a single noisy sample must not decide
  const int nf = s->fusable ? 2 : 1;
  float t[2][4] = {{1e30f, 1e30f, 1e30f, 1e30f}, {1e30f, 1e30f, 1e30f, 1e30f}};
  for (int pass = 0; pass < 2; pass++)
    for (int f = 0; f < nf; f++) {
      if (s->fusable) s->cc = f ? BX_CC_PLANE_FUSED : BX_CC_PLANE;
      for (int i = 0; i < 4; i++) {
        s->warps = cand[i];
        bool bad = launch(s, num_envs, 1, io, st) != 0;  // warm-up
        if (!bad) CUDA_OK(cudaEventRecord(e0, st));
        for (int r = 0; r < 8 && !bad; r++) bad = launch(s, num_envs, 1, io, st) != 0;
        if (bad) { s->warps = keep; s->cc = keep_cc; return 1; }
        CUDA_OK(cudaEventRecord(e1, st));
        CUDA_OK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
        t[f][i] = ms < t[f][i] ? ms : t[f][i];
      }
    }
  for (int f = 0; f < nf; f++)
    for (int i = 0; i < 4; i++)
      if (t[f][i] < best) {
        best = t[f][i];
        best_w = cand[i];
        best_cc = s->fusable ? (f ? BX_CC_PLANE_FUSED : BX_CC_PLANE) : keep_cc;
      }
  s->cc = best_cc;
  s->warps = best_w;
  s->warps_chosen = true;
  return 0;
}

extern "C" int bx_system_threads(const bx_system* s) { return s ? 32 * s->warps : -1; }

static int host_args_ok(const bx_system* s, int64_t B, const bx_step_io* hio) {
  if (!s || !hio) return fail("NULL argument");
  if (B <= 0) return fail("num_envs must be positive");
  if (hio->first_pos) return fail("auto-reset inputs are supported on the device path only (bx_step / bx_rollout)");
  if (hio->observe) return fail("observe is supported on the device path only (bx_step)");
  if (hio->policy) return fail("the in-kernel policy is supported on the device path only (bx_rollout)");
  if (hio->friction || hio->joint_offset)
    return fail("per-env friction / joint offsets are supported on the device path only (bx_step / bx_rollout)");
  if (!hio->pos || !hio->rot || !hio->vel || !hio->ang || !hio->act || !hio->pos_o || !hio->rot_o ||
      !hio->vel_o || !hio->ang_o)
    return fail("QP / act pointers must not be NULL");
  return 0;
}

// Zero-copy route of the host path: when every buffer is page-locked (bx_host_alloc, cudaHostRegister,
// torch pin_memory) the kernel works on the host memory directly -- each tile pulls its input spans
// over PCIe with bulk async copies and pushes its outputs with coalesced 128-bit stores, so upload,
// compute and download of different tiles overlap inside ONE launch on streams[slot].  *pinned = 0
// (nothing launched) when some buffer is pageable.  Does not synchronise.
// (Measured and rejected: outputs staged in device memory and moved by the copy engine, one DMA per
// contiguous host run -- 30-33 M env-steps/s for 8192 ants in 2-3 groups against 38 M for the tiles'
// own stores: the kernel -> copy serialisation costs more than the engine's higher link rate gains.)
static int host_launch_pinned(bx_system* s, int64_t B, const bx_step_io* hio, int slot, int* pinned_out) {
  const float* hin[BX_SEG_IN] = {hio->pos, hio->rot, hio->vel, hio->ang, hio->act};
  float* const* hout = io_outs(hio);
  const void* din[BX_SEG_IN];
  void* dout[BX_SEG_OUT];
  bool pinned = true;
  auto dev_ptr = [&](const void* hp, const void** dp) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, hp) != cudaSuccess) { cudaGetLastError(); return false; }
    if (at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) return false;
    *dp = at.devicePointer;
    return true;
  };
  for (int i = 0; i < BX_SEG_IN && pinned; i++) pinned = dev_ptr(hin[i], &din[i]);
  for (int i = 0; i < BX_SEG_OUT && pinned; i++) {
    dout[i] = nullptr;
    const bool constant = i >= 4 && s->h.seg_out[i].map < 0;
    if (hout[i] && s->h.seg_out[i].nper > 0 && !constant)
      pinned = dev_ptr(hout[i], const_cast<const void**>(&dout[i]));
  }
  const void* drand = nullptr;
  if (pinned && hio->env_rand) pinned = dev_ptr(hio->env_rand, &drand);
  *pinned_out = pinned ? 1 : 0;
  if (!pinned) return 0;
  bx_step_io d;
  memset(&d, 0, sizeof(d));
  d.env_rand = static_cast<const float*>(drand);
  d.pos = (const float*)din[0]; d.rot = (const float*)din[1]; d.vel = (const float*)din[2];
  d.ang = (const float*)din[3]; d.act = (const float*)din[4];
  for (int i = 0; i < BX_SEG_OUT; i++) io_outs(&d)[i] = static_cast<float*>(dout[i]);
  cudaStream_t st = s->streams[slot];
  // One wave of tiles (all resident together): stagger their starts over the time
  // the larger transfer direction needs at 60 GB/s (measured flat optimum 45-100,
  // profiles/r01_sweeps.txt), see bx_step_kernel.cuh.
  uint32_t stagger_ns = 0;
  if ((B + BX_ENVS_PER_CTA - 1) / BX_ENVS_PER_CTA <= 2 * static_cast<int64_t>(s->sm_count)) {
    int64_t h2d = 0, d2h = 0;
    bx_host_transfer_bytes(s, B, &h2d, &d2h);
    static const double gbps = getenv("BX_HOST_STAGGER_GBPS") ? atof(getenv("BX_HOST_STAGGER_GBPS")) : 60.0;
    if (gbps > 0) stagger_ns = static_cast<uint32_t>(static_cast<double>(std::max(h2d, d2h)) / gbps);
  }
  if (launch(s, B, 1, &d, st, stagger_ns)) return 1;
  for (int i = 4; i < BX_SEG_OUT; i++)
    if (hout[i] && s->h.seg_out[i].nper > 0 && s->h.seg_out[i].map < 0) {
      const float f = s->h.seg_out[i].fill;
      const size_t cnt = static_cast<size_t>(s->h.seg_out[i].nper) * B;
      if (f == 0.f) memset(hout[i], 0, cnt * 4); else std::fill(hout[i], hout[i] + cnt, f);
    }
  return 0;
}

static int host_streams(bx_system* s) {
  for (int i = 0; i < BX_HOST_STREAMS; i++)
    if (!s->streams[i]) CUDA_OK(cudaStreamCreateWithFlags(&s->streams[i], cudaStreamNonBlocking));
  return 0;
}

// The asynchronous form of the zero-copy host step: launches on internal stream `slot` and returns;
// bx_host_wait(sys, slot) waits for it.  Independent environment groups stepped through different
// slots overlap: one group's results travel to the host while another group's inputs arrive and a
// third computes (PCIe is full duplex; a single host-side stepping chain is latency bound).
extern "C" int bx_step_host_async(bx_system* s, int64_t B, const bx_step_io* hio, int32_t slot) {
  if (host_args_ok(s, B, hio)) return 1;
  if (slot < 0 || slot >= BX_HOST_STREAMS) return fail("slot must be 0, 1 or 2");
  CUDA_OK(cudaSetDevice(s->device));
  if (host_streams(s)) return 1;
  int pinned = 0;
  if (host_launch_pinned(s, B, hio, slot, &pinned)) return 1;
  if (!pinned) return fail("bx_step_host_async needs page-locked buffers (bx_host_alloc / cudaHostRegister / pin_memory); "
                           "pageable buffers go through bx_step_host");
  return 0;
}
extern "C" int bx_host_wait(bx_system* s, int32_t slot) {
  if (!s) return fail("system is NULL");
  if (slot < 0 || slot >= BX_HOST_STREAMS) return fail("slot must be 0, 1 or 2");
  if (s->streams[slot]) CUDA_OK(cudaStreamSynchronize(s->streams[slot]));
  return 0;
}

// Host-buffer step: H2D copies, one launch, D2H copies, synchronise.
extern "C" int bx_step_host(bx_system* s, int64_t B, const bx_step_io* hio) {
  if (host_args_ok(s, B, hio)) return 1;
  CUDA_OK(cudaSetDevice(s->device));
  const int64_t n = s->h.nb, a = s->h.act_size;
  // in: 13n + a ; out: the segments' nper (floats per env)
  if (host_streams(s)) return 1;
  {
    int pinned = 0;
    if (host_launch_pinned(s, B, hio, 0, &pinned)) return 1;
    if (pinned) {
      CUDA_OK(cudaStreamSynchronize(s->streams[0]));
      return 0;
    }
  }
  size_t per_env = static_cast<size_t>(13 * n + a);
  for (int i = 0; i < BX_SEG_OUT; i++) per_env += s->h.seg_out[i].nper;
  if (hio->env_rand) per_env += 3;
  const size_t need = per_env * B;
  if (need > s->scratch_floats) {
    cudaFree(s->d_scratch);
    s->d_scratch = nullptr;
    s->scratch_floats = 0;
    CUDA_OK(cudaMalloc(&s->d_scratch, need * 4));
    s->scratch_floats = need;
  }
  float* cur = s->d_scratch;
  auto take = [&](size_t per) { float* r = cur; cur += per * B; return r; };
  // device arrays, in the order of the kernel's segments
  const size_t in_per[BX_SEG_IN] = {(size_t)3 * n, (size_t)4 * n, (size_t)3 * n, (size_t)3 * n, (size_t)a};
  const float* h_in[BX_SEG_IN] = {hio->pos, hio->rot, hio->vel, hio->ang, hio->act};
  float* d_in[BX_SEG_IN];
  for (int i = 0; i < BX_SEG_IN; i++) d_in[i] = take(in_per[i]);
  size_t out_per[BX_SEG_OUT];
  for (int i = 0; i < BX_SEG_OUT; i++) out_per[i] = static_cast<size_t>(s->h.seg_out[i].nper);
  float* const* h_out = io_outs(hio);
  float* d_out[BX_SEG_OUT];
  for (int i = 0; i < BX_SEG_OUT; i++) d_out[i] = take(out_per[i]);
  float* d_rand = hio->env_rand ? take(3) : nullptr;
  // Pipeline: the batch is cut into chunks (multiples of the 32-env tile); chunk
  // c's H2D, kernel and D2H run on stream c % 3, so the upload of one chunk, the
  // step of another and the download of a third overlap (PCIe is full duplex).
  // Outputs that are constants of the dynamics mode (info.joint / info.actuator.vel
  // in pbd) are not transferred: the host fills them while the GPU works.
  int64_t chunk = ((B + 3) / 4 + 31) / 32 * 32;
  if (chunk < 1024) chunk = std::min<int64_t>(B, 1024);
  int ci = 0;
  for (int64_t e0 = 0; e0 < B; e0 += chunk, ci++) {
    const int64_t nb_ = std::min<int64_t>(chunk, B - e0);
    cudaStream_t st = s->streams[ci % BX_HOST_STREAMS];
    bx_step_io d;
    memset(&d, 0, sizeof(d));
    for (int i = 0; i < BX_SEG_IN; i++)
      CUDA_OK(cudaMemcpyAsync(d_in[i] + e0 * in_per[i], h_in[i] + e0 * in_per[i], in_per[i] * nb_ * 4,
                              cudaMemcpyHostToDevice, st));
    d.pos = d_in[0] + e0 * in_per[0]; d.rot = d_in[1] + e0 * in_per[1];
    d.vel = d_in[2] + e0 * in_per[2]; d.ang = d_in[3] + e0 * in_per[3];
    d.act = d_in[4] + e0 * in_per[4];
    if (d_rand) {
      CUDA_OK(cudaMemcpyAsync(d_rand + e0 * 3, hio->env_rand + e0 * 3, 3 * nb_ * 4, cudaMemcpyHostToDevice, st));
      d.env_rand = d_rand + e0 * 3;
    }
    float** dp = io_outs(&d);
    for (int i = 0; i < BX_SEG_OUT; i++) {
      const bool constant = s->h.seg_out[i].map < 0;
      dp[i] = (i < 4 || (h_out[i] && out_per[i] && !constant)) ? d_out[i] + e0 * out_per[i] : nullptr;
    }
    if (launch(s, nb_, 1, &d, st)) return 1;
    for (int i = 0; i < BX_SEG_OUT; i++)
      if (dp[i] && h_out[i])
        CUDA_OK(cudaMemcpyAsync(h_out[i] + e0 * out_per[i], dp[i], out_per[i] * nb_ * 4,
                                cudaMemcpyDeviceToHost, st));
  }
  for (int i = 4; i < BX_SEG_OUT; i++)
    if (h_out[i] && out_per[i] && s->h.seg_out[i].map < 0) {
      const float f = s->h.seg_out[i].fill;
      if (f == 0.f) memset(h_out[i], 0, out_per[i] * B * 4);
      else std::fill(h_out[i], h_out[i] + out_per[i] * B, f);
    }
  for (int i = 0; i < BX_HOST_STREAMS; i++) CUDA_OK(cudaStreamSynchronize(s->streams[i]));
  return 0;
}

// Bytes bx_step_host moves over PCIe for `num_envs` environments (constant
// outputs are filled on the host and not counted).
extern "C" int bx_host_transfer_bytes(const bx_system* s, int64_t B, int64_t* h2d, int64_t* d2h) {
  if (!s) return fail("NULL argument");
  const int64_t n = s->h.nb, a = s->h.act_size;
  if (h2d) *h2d = 4 * B * (13 * n + a);
  int64_t o = 0;
  for (int i = 0; i < 13; i++)  // the step's own outputs; the optional angle_vel epilogue adds 24 J bytes/env
    if (i < 4 || s->h.seg_out[i].map >= 0) o += s->h.seg_out[i].nper;
  if (d2h) *d2h = 4 * B * o;
  return 0;
}

// Page-locked host memory for bx_step_host buffers (pageable memory works too,
// at a fraction of the PCIe rate).
extern "C" void* bx_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
    fail("cudaHostAlloc failed");
    return nullptr;
  }
  return p;
}
extern "C" void bx_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- bx_step_host_static: the host-buffer step for callers that may not allocate ----
// (no counterpart in the reference, see
// include/brax_b200.h).  Same contract as bx_step_host's staged route -- copy in, ONE
// launch, copy out, synchronise -- but the device staging area is this file-scope array
// instead of a cudaMalloc'ed buffer, the work runs on the CALLER's stream (no stream is
// created) and nothing is allocated or freed.  Every GPU that loads the library holds the
// array (BX_STATIC_SCRATCH_BYTES) until exit.
//
// Contents: the array keeps whatever the previous call left.  Nothing in it has to start
// at zero: the input part is overwritten by this call's H2D copies before the launch and
// the kernel writes every requested output in full (zero_info is done in shared memory,
// system.py:76-84), so stale contents are never read.  Calls that overlap on the GPU
// (two host threads, two streams) share the array and RACE: one call at a time per device.
static __device__ __align__(256) float bx_static_scratch[BX_STATIC_SCRATCH_BYTES / 4];

#define BX_STATIC_MAX_DEVICES 64
static float* g_static_ptr[BX_STATIC_MAX_DEVICES];  // device address of the array, per device
static std::atomic<int> g_static_state[BX_STATIC_MAX_DEVICES];  // 0 = not looked up, 2 = ready

// The array's NAME is a host-side stand-in in host code, not a device address: look the
// address up once per device, with that device current.
static int static_scratch_ptr(int device, float** out) {
  if (device < 0 || device >= BX_STATIC_MAX_DEVICES) return fail("bx_step_host_static: device index out of range");
  if (g_static_state[device].load(std::memory_order_acquire) != 2) {
    void* p = nullptr;
    CUDA_OK(cudaGetSymbolAddress(&p, bx_static_scratch));  // s->device is current (caller)
    if (!p) return fail("cudaGetSymbolAddress returned NULL");
    g_static_ptr[device] = static_cast<float*>(p);  // every thread that gets here stores the same value
    g_static_state[device].store(2, std::memory_order_release);
  }
  *out = g_static_ptr[device];
  return 0;
}

extern "C" int64_t bx_static_scratch_bytes(void) { return BX_STATIC_SCRATCH_BYTES; }

extern "C" int bx_step_host_static(const bx_system* s, int64_t B, const bx_step_io* hio, void* stream) {
  if (!s || !hio) return fail("NULL argument");
  if (B <= 0) return fail("num_envs must be positive");
  if (hio->first_pos) return fail("auto-reset inputs are supported on the device path only (bx_step / bx_rollout)");
  if (hio->observe) return fail("observe is supported on the device path only (bx_step)");
  if (hio->policy) return fail("the in-kernel policy is supported on the device path only (bx_rollout)");
  if (hio->friction || hio->joint_offset)
    return fail("per-env friction / joint offsets are supported on the device path only (bx_step / bx_rollout)");
  if (!hio->pos || !hio->rot || !hio->vel || !hio->ang || !hio->act || !hio->pos_o || !hio->rot_o ||
      !hio->vel_o || !hio->ang_o)
    return fail("QP / act pointers must not be NULL");
  const int64_t n = s->h.nb, a = s->h.act_size;
  float* const* h_out = io_outs(hio);
  // layout of the staging area: inputs, then the requested non-constant outputs, each
  // sub-array on a 256-byte boundary (the kernel's 128-bit / bulk paths need 16)
  const size_t in_per[BX_SEG_IN] = {(size_t)3 * n, (size_t)4 * n, (size_t)3 * n, (size_t)3 * n, (size_t)a};
  size_t out_per[BX_SEG_OUT];
  bool staged[BX_SEG_OUT];
  auto round64 = [](size_t floats) { return (floats + 63) / 64 * 64; };
  size_t need = 0;
  for (int i = 0; i < BX_SEG_IN; i++) need += round64(in_per[i] * B);
  for (int i = 0; i < BX_SEG_OUT; i++) {
    out_per[i] = static_cast<size_t>(s->h.seg_out[i].nper);
    staged[i] = i < 4 || (h_out[i] && out_per[i] && s->h.seg_out[i].map >= 0);
    if (staged[i]) need += round64(out_per[i] * B);
  }
  if (hio->env_rand) need += round64((size_t)3 * B);
  if (need * 4 > static_cast<size_t>(BX_STATIC_SCRATCH_BYTES)) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "bx_step_host_static: %lld environments need %zu bytes of staging, the static scratch array holds "
             "%lld (maximum %lld environments of this system with these outputs); use bx_step_host",
             static_cast<long long>(B), need * 4, static_cast<long long>(BX_STATIC_SCRATCH_BYTES),
             static_cast<long long>(static_cast<size_t>(BX_STATIC_SCRATCH_BYTES) / 4 / (need / B + 1)));
    return fail(msg);
  }
  int prev_device = -1;
  CUDA_OK(cudaGetDevice(&prev_device));
  if (prev_device != s->device) CUDA_OK(cudaSetDevice(s->device));
  struct Restore {
    int prev, cur;
    ~Restore() { if (prev != cur) cudaSetDevice(prev); }
  } restore{prev_device, s->device};
  float* cur = nullptr;
  if (static_scratch_ptr(s->device, &cur)) return 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto take = [&](size_t floats) { float* r = cur; cur += round64(floats); return r; };
  const float* h_in[BX_SEG_IN] = {hio->pos, hio->rot, hio->vel, hio->ang, hio->act};
  float* d_in[BX_SEG_IN];
  for (int i = 0; i < BX_SEG_IN; i++) {
    d_in[i] = take(in_per[i] * B);
    CUDA_OK(cudaMemcpyAsync(d_in[i], h_in[i], in_per[i] * B * 4, cudaMemcpyHostToDevice, st));
  }
  bx_step_io d;
  memset(&d, 0, sizeof(d));
  d.pos = d_in[0]; d.rot = d_in[1]; d.vel = d_in[2]; d.ang = d_in[3]; d.act = d_in[4];
  float** dp = io_outs(&d);
  for (int i = 0; i < BX_SEG_OUT; i++) dp[i] = staged[i] ? take(out_per[i] * B) : nullptr;
  if (hio->env_rand) {
    float* d_rand = take((size_t)3 * B);
    CUDA_OK(cudaMemcpyAsync(d_rand, hio->env_rand, (size_t)3 * B * 4, cudaMemcpyHostToDevice, st));
    d.env_rand = d_rand;
  }
  if (launch(s, B, 1, &d, st)) return 1;
  for (int i = 0; i < BX_SEG_OUT; i++)
    if (dp[i] && h_out[i])
      CUDA_OK(cudaMemcpyAsync(h_out[i], dp[i], out_per[i] * B * 4, cudaMemcpyDeviceToHost, st));
  // outputs that are constants of the dynamics mode: filled here, not transferred
  for (int i = 4; i < BX_SEG_OUT; i++)
    if (h_out[i] && out_per[i] && s->h.seg_out[i].map < 0) {
      const float f = s->h.seg_out[i].fill;
      if (f == 0.f) memset(h_out[i], 0, out_per[i] * B * 4);
      else std::fill(h_out[i], h_out[i] + out_per[i] * B, f);
    }
  CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}
