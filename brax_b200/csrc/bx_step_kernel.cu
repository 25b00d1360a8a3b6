// bx_step_kernel.cu -- the fused Brax v1 `System.step` kernel for sm_100a.
//
// One launch = one (or T) System.step for all environments
// (reference: brax/physics/system.py:244-377; XLA runs this as a few hundred
// small fusions per substep with the state round-tripping through memory).
//
// Mapping (DESIGN.md "kernel"):
//   * a CTA owns a tile of 32 environments; lane l of EVERY warp works on
//     environment l of the tile, so all control flow and all table indices are
//     warp-uniform (no divergence, table reads are shared-memory broadcasts);
//   * the W warps of the CTA split the joints / bodies / contact rows of those
//     32 environments between them (warp w takes items w, w+W, ...), phases are
//     separated by __syncthreads();
//   * the whole QP state of the tile, the per-joint / per-contact deltas and the
//     Info accumulators live in shared memory for the entire substep loop, laid
//     out [row][env] with a 33-float row stride: compute accesses (row*33+lane)
//     and the transposing global I/O (consecutive rows, same env) are both
//     bank-conflict free;
//   * HBM is touched once on the way in and once on the way out, with
//     coalesced 128-bit loads/stores of the boundary layout [B,N,3|4].
// FP32 CUDA-core bound by design; tensor cores do not apply (no contraction).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/brax_b200.h"
#include "bx_math.cuh"
#include "bx_tables.h"

struct BxKernelParams {
  BxHeader h;
  const int32_t* tab_i;
  const float* tab_f;
  bx_step_io io;
  int64_t num_envs;
  int32_t num_steps;  // >1: rollout, act is [T,B,A]
};

#define RS BX_ROW_STRIDE

struct Ctx {
  float* st;         // shared state base + lane
  const float* tf;   // float tables (shared)
  const int32_t* ti; // int tables (shared)
  int warp, nwarp, lane;
  int cur;           // ping-pong index of the current pos/rot buffer
};

#define ROW(c, r) ((c).st[(r) * RS])

BXD V3 ld3(const Ctx& c, int r) { return v3(ROW(c, r), ROW(c, r + 1), ROW(c, r + 2)); }
BXD Q4 ld4(const Ctx& c, int r) {
  return q4(ROW(c, r), ROW(c, r + 1), ROW(c, r + 2), ROW(c, r + 3));
}
BXD void st3(const Ctx& c, int r, V3 v) { ROW(c, r) = v.x; ROW(c, r + 1) = v.y; ROW(c, r + 2) = v.z; }
BXD void st4(const Ctx& c, int r, Q4 q) {
  ROW(c, r) = q.w; ROW(c, r + 1) = q.x; ROW(c, r + 2) = q.y; ROW(c, r + 3) = q.z;
}
BXD V3 tf3(const float* t) { return v3(t[0], t[1], t[2]); }

// ---- global <-> shared transposing I/O ---------------------------------------
// g: [B, nper] row-major; this CTA owns envs [env0, env0+nenv).  Element
// (e, r) goes to shared row (row0 + r), env slot e.
__device__ void load_rows(float* sm, const float* __restrict__ g, int64_t env0,
                          int nenv, int nper, int row0) {
  const float* src = g + env0 * nper;
  int total = nenv * nper;
  if ((total & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    for (int i = threadIdx.x; i < (total >> 2); i += blockDim.x) {
      float4 v = __ldg(s4 + i);
      float vv[4] = {v.x, v.y, v.z, v.w};
      int idx = i << 2;
      int e = idx / nper, r = idx - e * nper;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        sm[(row0 + r) * RS + e] = vv[k];
        if (++r == nper) { r = 0; e++; }
      }
    }
  } else {
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
      int e = idx / nper, r = idx - e * nper;
      sm[(row0 + r) * RS + e] = __ldg(src + idx);
    }
  }
}

// map == nullptr: element (e, r) comes from row (row0 + r); otherwise from row
// map[r], or 0 when map[r] < 0.
__device__ void store_rows(const float* sm, float* __restrict__ g, int64_t env0,
                           int nenv, int nper, int row0, const int32_t* map) {
  if (g == nullptr) return;
  float* dst = g + env0 * nper;
  int total = nenv * nper;
  if ((total & 3) == 0 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int i = threadIdx.x; i < (total >> 2); i += blockDim.x) {
      float vv[4];
      int idx = i << 2;
      int e = idx / nper, r = idx - e * nper;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        int row = map ? map[r] : row0 + r;
        vv[k] = row >= 0 ? sm[row * RS + e] : 0.f;
        if (++r == nper) { r = 0; e++; }
      }
      d4[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
    }
  } else {
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
      int e = idx / nper, r = idx - e * nper;
      int row = map ? map[r] : row0 + r;
      dst[idx] = row >= 0 ? sm[row * RS + e] : 0.f;
    }
  }
}

__device__ void store_const(float* __restrict__ g, int64_t env0, int nenv, int nper,
                            float value) {
  if (g == nullptr) return;
  float* dst = g + env0 * nper;
  int total = nenv * nper;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) dst[idx] = value;
}

// ---- joints: axis / angle -------------------------------------------------
// Revolute.axis_angle joints.py:311-319 (spring_joints.py:150-157)
BXD void axis_angle_revolute(const float* jf, Q4 rp, Q4 rc, V3* axis, float* angle) {
  V3 ax = rotate(tf3(jf + JF_AXIS_P), rp);
  V3 ref_p = rotate(tf3(jf + JF_AXIS_P + 6), rp);
  V3 ref_c = rotate(tf3(jf + JF_AXIS_C + 6), rc);
  axis[0] = ax;
  angle[0] = signed_angle(ax, ref_p, ref_c);
}

// Spherical.axis_angle joints.py:388-415 (spring_joints.py:212-224, 276-303)
BXD void axis_angle_spherical(const float* jf, Q4 rp, Q4 rc, V3* axis, float* angle) {
  V3 a1p = rotate(tf3(jf + JF_AXIS_P), rp);
  V3 a2p = rotate(tf3(jf + JF_AXIS_P + 3), rp);
  V3 a1c = rotate(tf3(jf + JF_AXIS_C), rc);
  V3 a2c = rotate(tf3(jf + JF_AXIS_C + 3), rc);
  V3 a3c = rotate(tf3(jf + JF_AXIS_C + 6), rc);
  V3 lon = cross(a3c, a1p);
  lon = vdiv(lon, 1e-10f + safe_norm(lon));
  float psi = signed_angle(a1p, a2p, lon);
  V3 xz = dot(a1p, a1c) * a1c + dot(a1p, a2c) * a2c;
  xz = vdiv(xz, 1e-10f + safe_norm(xz));
  float theta = acosf(clipf(dot(xz, a1p), -1.f, 1.f)) * signf(dot(a1p, a3c));
  float phi = signed_angle(-a3c, a2c, lon);
  axis[0] = a1p; axis[1] = a2c; axis[2] = a3c;
  angle[0] = psi; angle[1] = theta; angle[2] = phi;
}

BXD int joint_dof(int kind) {
  return (kind == BX_JOINT_REVOLUTE || kind == BX_JOINT_SPRING_REVOLUTE)
             ? 1
             : (kind == BX_JOINT_SPRING_UNIVERSAL ? 2 : 3);
}

BXD void joint_axis_angle(int kind, const float* jf, Q4 rp, Q4 rc, V3* axis,
                          float* angle) {
  if (kind == BX_JOINT_REVOLUTE || kind == BX_JOINT_SPRING_REVOLUTE)
    axis_angle_revolute(jf, rp, rc, axis, angle);
  else
    axis_angle_spherical(jf, rp, rc, axis, angle);
}

// ---- phase T: actuator torques and joint damping torques -------------------
// Writes the torque VECTOR of actuator a to rows JS+3a and the damping torque of
// joint j to rows JS+3(na+j); the body phase turns them into angular
// accelerations (actuators.py:52-112, joints.py:103-128).
__device__ __forceinline__ void phase_torques(const Ctx& c, const BxHeader& h, bool damp) {
  const int rrot = c.cur ? h.r_rot1 : h.r_rot0;
  for (int a = c.warp; a < h.na; a += c.nwarp) {
    const int32_t* ai = c.ti + h.io_act + a * AI_STRIDE;
    const int j = ai[AI_JOINT];
    const int32_t* ji = c.ti + h.io_joint + j * JI_STRIDE;
    const float* jf = c.tf + h.fo_joint + j * JF_STRIDE;
    const int kind = ji[JI_KIND];
    Q4 rp = ld4(c, rrot + 4 * ji[JI_BODY_P]);
    Q4 rc = ld4(c, rrot + 4 * ji[JI_BODY_C]);
    V3 axis[3];
    float angle[3];
    joint_axis_angle(kind, jf, rp, rc, axis, angle);
    const float strength = c.tf[h.fo_act + a];
    const int dof = joint_dof(kind);
    V3 tq = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      if (d < dof) {
        int idx = ai[AI_IDX + d];
        // take(act, idx, mode=clip) * act_mask  (actuators.py:64)
        float m = idx >= 0 ? 1.f : 0.f;
        idx = min(max(idx, 0), h.act_size - 1);
        float av = ROW(c, h.r_act + idx) * m;
        float lo = jf[JF_LIMIT + 2 * d], hi = jf[JF_LIMIT + 2 * d + 1];
        float t;
        if (ai[AI_KIND] == BX_ACT_TORQUE) {
          t = av * strength * -1.f;
          t = angle[d] < lo ? 0.f : t;
          t = angle[d] > hi ? 0.f : t;
        } else {
          float tgt = clipf(av * 3.14159265358979323846f / 180.f, lo, hi);
          t = (tgt - angle[d]) * strength;
        }
        tq = tq + axis[d] * t;
      }
    }
    st3(c, h.r_js + 3 * a, tq);
  }
  if (damp) {
    for (int j = c.warp; j < h.nj; j += c.nwarp) {
      const int32_t* ji = c.ti + h.io_joint + j * JI_STRIDE;
      const float* jf = c.tf + h.fo_joint + j * JF_STRIDE;
      V3 ap = ld3(c, h.r_ang + 3 * ji[JI_BODY_P]);
      V3 ac = ld3(c, h.r_ang + 3 * ji[JI_BODY_C]);
      V3 tq = (-1.f * jf[JF_ANG_DAMP]) * (ap - ac);
      st3(c, h.r_js + 3 * (h.na + j), tq);
    }
  }
}

// Sum of the angular accelerations of body b from actuators (dp_a), joint
// damping (dp_j) and forces (dp_f); system.py:268-271.
BXD void body_accels(const Ctx& c, const BxHeader& h, int b, const int32_t* bi,
                     const float* bf, bool damp, V3* dpa_ang, V3* dpj_ang,
                     V3* dpf_vel, V3* dpf_ang) {
  V3 inv_i = tf3(bf + BF_INV_INERTIA);
  V3 dpa = v3(0.f, 0.f, 0.f), dpj = v3(0.f, 0.f, 0.f);
  {
    const int32_t* adj = c.ti + h.io_adj + bi[BI_AADJ_OFF];
    const int n = bi[BI_AADJ_N];
    for (int i = 0; i < n; i++) {
      int e = adj[i];
      V3 t = inv_i * ld3(c, h.r_js + 3 * (e >> 2));
      // Torque: parent +, child -; Angle: parent -, child + (actuators.py:88-89,109-110)
      dpa = (e & 1) ? dpa - t : dpa + t;
    }
  }
  if (damp) {
    const int32_t* adj = c.ti + h.io_adj + bi[BI_JADJ_OFF];
    const int n = bi[BI_JADJ_N];
    for (int i = 0; i < n; i++) {
      int e = adj[i];
      V3 t = inv_i * ld3(c, h.r_js + 3 * (h.na + (e >> 1)));
      dpj = (e & 1) ? dpj - t : dpj + t;
    }
  }
  V3 fv = v3(0.f, 0.f, 0.f), fa = v3(0.f, 0.f, 0.f);
  {
    const int32_t* adj = c.ti + h.io_adj + bi[BI_FADJ_OFF];
    const int n = bi[BI_FADJ_N];
    for (int i = 0; i < n; i++) {
      int f = adj[i];
      const int32_t* fi = c.ti + h.io_force + f * FI_STRIDE;
      float strength = c.tf[h.fo_force + f];
      V3 a = v3(ROW(c, h.r_act + min(max(fi[FI_IDX], 0), h.act_size - 1)),
                ROW(c, h.r_act + min(max(fi[FI_IDX + 1], 0), h.act_size - 1)),
                ROW(c, h.r_act + min(max(fi[FI_IDX + 2], 0), h.act_size - 1)));
      V3 d = vdiv(a * strength, bf[BF_MASS]);  // forces.py:44,87
      if (fi[FI_KIND] == BX_FORCE_THRUSTER) fv = fv + d; else fa = fa + d;
    }
  }
  *dpa_ang = dpa; *dpj_ang = dpj; *dpf_vel = fv; *dpf_ang = fa;
}

// ---- integrator pieces (per body) -------------------------------------------
// Euler.kinetic integrators.py:50-68
BXD void kinetic(const BxHeader& h, const float* bf, V3 vel, V3 ang, V3* pos, Q4* rot) {
  V3 pm = tf3(bf + BF_POS_MASK), rm = tf3(bf + BF_ROT_MASK);
  *pos = *pos + (vel * h.dt) * pm;
  V3 w = (ang * rm) * h.half_dt;
  Q4 r = *rot + quat_mul(q4(0.f, w.x, w.y, w.z), *rot);
  float n = bx_sqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
  float rn = bx_rcp(n);
  *rot = q4(bx_mul_rcp(r.w, n, rn), bx_mul_rcp(r.x, n, rn), bx_mul_rcp(r.y, n, rn), bx_mul_rcp(r.z, n, rn));
}

// Euler.velocity_projection integrators.py:122-146
BXD void velocity_projection(const BxHeader& h, const float* bf, V3 pos, V3 ppos,
                             Q4 prot, Q4* rot, V3* vel, V3* ang) {
  V3 pm = tf3(bf + BF_POS_MASK), rm = tf3(bf + BF_ROT_MASK);
  Q4 r = *rot;
  float n = bx_sqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
  float rn = bx_rcp(n);
  Q4 nr = q4(bx_mul_rcp(r.w, n, rn), bx_mul_rcp(r.x, n, rn), bx_mul_rcp(r.y, n, rn), bx_mul_rcp(r.z, n, rn));
  *vel = ((pos - ppos) * h.inv_dt) * pm;
  Q4 dq = quat_mul(nr, qconj(prot));
  float sg = dq.w >= 0.f ? 1.f : -1.f;
  V3 a = (2.f * v3(dq.x, dq.y, dq.z)) * h.inv_dt;
  V3 scale = sg * rm;
  *ang = (scale * a) * rm;
  *rot = nr;
}

// ---- phase B1 (pbd): acceleration update + kinetic step, per body ------------
// system.py:268-274 / :287-293.  Reads pos/rot from buffer cur, writes the
// advanced pos/rot into buffer cur^1 (buffer cur then IS qprev).
__device__ __forceinline__ void phase_body_acc_kinetic(const Ctx& c, const BxHeader& h,
                                                    bool accumulate_info) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  const int npos = c.cur ? h.r_pos0 : h.r_pos1, nrot = c.cur ? h.r_rot0 : h.r_rot1;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    V3 dpa, dpj, fv, fa;
    body_accels(c, h, b, bi, bf, true, &dpa, &dpj, &fv, &fa);
    V3 acc_ang = (dpa + fa) + dpj;
    V3 vel = ld3(c, h.r_vel + 3 * b), ang = ld3(c, h.r_ang + 3 * b);
    V3 g = v3(h.gravity[0], h.gravity[1], h.gravity[2]);
    // Euler.update acc branch integrators.py:87-95
    vel = (h.vel_damp * vel + (fv + g) * h.dt) * tf3(bf + BF_POS_MASK);
    ang = (h.ang_damp * ang + acc_ang * h.dt) * tf3(bf + BF_ROT_MASK);
    V3 pos = ld3(c, rpos + 3 * b);
    Q4 rot = ld4(c, rrot + 4 * b);
    kinetic(h, bf, vel, ang, &pos, &rot);
    st3(c, h.r_vel + 3 * b, vel);
    st3(c, h.r_ang + 3 * b, ang);
    st3(c, npos + 3 * b, pos);
    st4(c, nrot + 4 * b, rot);
    if (accumulate_info && bi[BI_ASLOT] >= 0) {  // info.actuator += dp_a (system.py:315)
      int r = h.r_ia + 3 * bi[BI_ASLOT];
      st3(c, r, ld3(c, r) + dpa);
    }
  }
}

// ---- pbd joints: joints.py ---------------------------------------------------
struct DQ { V3 pos; Q4 rot; };

// Joint.apply_angle_update joints.py:130-152 (rot parts only; pos parts are 0)
BXD void angle_update(V3 dq, V3 ii_p, V3 ii_c, Q4 rp, Q4 rc, float scale_ang,
                      Q4* out_p, Q4* out_c) {
  float th = safe_norm(dq);
  V3 n = vdiv(dq, th + 1e-6f);
  float w1 = dot(n, ii_p * n), w2 = dot(n, ii_c * n);
  float dl = bx_div(-th, w1 + w2 + 1e-6f);
  V3 p = (-dl) * n;
  *out_p = scale_ang * (.5f * vec_quat_mul(ii_p * p, rp));
  *out_c = scale_ang * (-.5f * vec_quat_mul(ii_c * p, rc));
}

// Joint.apply_position_update joints.py:154-195
BXD void position_update(V3 wp, V3 wc, V3 pos_p, V3 pos_c, Q4 rp, Q4 rc,
                         float im_p, float im_c, float m_p, float m_c, V3 ii_p,
                         V3 ii_c, float scale_pos, DQ* dp, DQ* dc) {
  V3 dx = wp - wc;
  V3 r_p = wp - pos_p, r_c = wc - pos_c;
  float cn = safe_norm(dx);
  V3 n = vdiv(dx, cn + 1e-6f);
  V3 cr1 = cross(r_p, n), cr2 = cross(r_c, n);
  float w1 = im_p + dot(cr1, ii_p * cr1);
  float w2 = im_c + dot(cr2, ii_c * cr2);
  float dl = bx_div(-cn, w1 + w2 + 1e-6f);
  V3 p = dl * n;
  dp->pos = scale_pos * vdiv(p, m_p);
  dp->rot = scale_pos * (.5f * vec_quat_mul(ii_p * cross(r_p, p), rp));
  dc->pos = scale_pos * vdiv(-p, m_c);
  dc->rot = scale_pos * (-.5f * vec_quat_mul(ii_c * cross(r_c, p), rc));
}

// limit_angle closure of Spherical.apply_reduced joints.py:343-354
BXD void limit_angle(V3 n, V3 n_1, V3 n_2, float lo, float hi, V3 ii_p, V3 ii_c,
                     Q4 rp, Q4 rc, float scale_ang, Q4* out_p, Q4* out_c) {
  float ph = signed_angle(n, n_1, n_2);
  float mask = ph < lo ? 1.f : 0.f;
  mask = ph > hi ? 1.f : mask;
  ph = clipf(ph, lo, hi);
  V3 n1 = rotate(n_1, quat_rot_axis(n, ph));
  V3 dq = cross(n1, n_2) * mask;
  angle_update(dq, ii_p, ii_c, rp, rc, scale_ang, out_p, out_c);
}

// ---- phase J (pbd): Joint.apply for every joint -> per-joint delta slots -------
// joints.py:79-100, Revolute :270-309, Spherical :332-386.  Slot of joint j:
// rows JS+14j .. +6 = parent (pos3, rot4), +7 .. +13 = child.
__device__ __forceinline__ void phase_joints_pbd(const Ctx& c, const BxHeader& h) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  for (int j = c.warp; j < h.nj; j += c.nwarp) {
    const int32_t* ji = c.ti + h.io_joint + j * JI_STRIDE;
    const float* jf = c.tf + h.fo_joint + j * JF_STRIDE;
    const int bp = ji[JI_BODY_P], bc = ji[JI_BODY_C];
    const float* bfp = c.tf + h.fo_body + bp * BF_STRIDE;
    const float* bfc = c.tf + h.fo_body + bc * BF_STRIDE;
    V3 pos_p = ld3(c, rpos + 3 * bp), pos_c = ld3(c, rpos + 3 * bc);
    Q4 rp = ld4(c, rrot + 4 * bp), rc = ld4(c, rrot + 4 * bc);
    V3 ii_p = tf3(bfp + BF_INV_INERTIA), ii_c = tf3(bfc + BF_INV_INERTIA);
    V3 wp = pos_p + rotate(tf3(jf + JF_OFF_P), rp);
    V3 wc = pos_c + rotate(tf3(jf + JF_OFF_C), rc);
    DQ dp, dc;
    position_update(wp, wc, pos_p, pos_c, rp, rc, bfp[BF_INV_MASS], bfc[BF_INV_MASS],
                    bfp[BF_MASS], bfc[BF_MASS], ii_p, ii_c, jf[JF_SCALE_POS], &dp, &dc);
    const float sa = jf[JF_SCALE_ANG];
    if (ji[JI_KIND] == BX_JOINT_REVOLUTE) {
      V3 axis = rotate(tf3(jf + JF_AXIS_P), rp);
      V3 ref_p = rotate(tf3(jf + JF_AXIS_P + 6), rp);
      V3 ref_c = rotate(tf3(jf + JF_AXIS_C + 6), rc);
      float psi = signed_angle(axis, ref_p, ref_c);
      V3 axis_c = rotate(tf3(jf + JF_AXIS_C), rc);
      V3 dq_1 = cross(axis, axis_c);
      float ph = clipf(psi, jf[JF_LIMIT], jf[JF_LIMIT + 1]);
      V3 n1 = rotate(ref_p, quat_rot_axis(axis, ph));
      V3 dq_2 = cross(n1, ref_c);
      Q4 p1, c1, p2, c2;
      angle_update(dq_1, ii_p, ii_c, rp, rc, sa, &p1, &c1);
      angle_update(dq_2, ii_p, ii_c, rp, rc, sa, &p2, &c2);
      dp.rot = dp.rot + (p1 + p2);
      dc.rot = dc.rot + (c1 + c2);
    } else {
      V3 a1p = rotate(tf3(jf + JF_AXIS_P), rp);
      V3 a2p = rotate(tf3(jf + JF_AXIS_P + 3), rp);
      V3 a1c = rotate(tf3(jf + JF_AXIS_C), rc);
      V3 a2c = rotate(tf3(jf + JF_AXIS_C + 3), rc);
      V3 a3c = rotate(tf3(jf + JF_AXIS_C + 6), rc);
      V3 lon = cross(a3c, a1p);
      lon = vdiv(lon, 1e-6f + safe_norm(lon));
      V3 xz = dot(a1p, a1c) * a1c + dot(a1p, a2c) * a2c;
      xz = vdiv(xz, 1e-6f + safe_norm(xz));
      V3 a2n = cross(xz, a1p);
      a2n = vdiv(a2n, 1e-6f + safe_norm(a2n));
      float sg = signf(dot(a1p, a3c));
      Q4 p1, c1, p2, c2, p3, c3;
      limit_angle(a1p, a2p, lon, jf[JF_LIMIT], jf[JF_LIMIT + 1], ii_p, ii_c, rp, rc, sa, &p1, &c1);
      limit_angle((-a2n) * sg, a1p, xz, jf[JF_LIMIT + 2], jf[JF_LIMIT + 3], ii_p, ii_c, rp, rc, sa, &p2, &c2);
      limit_angle(a3c, lon, a2c, jf[JF_LIMIT + 4], jf[JF_LIMIT + 5], ii_p, ii_c, rp, rc, sa, &p3, &c3);
      dp.rot = dp.rot + ((p1 + p2) + p3);
      dc.rot = dc.rot + ((c1 + c2) + c3);
    }
    const int r = h.r_js + 14 * j;
    st3(c, r, dp.pos); st4(c, r + 3, dp.rot);
    st3(c, r + 7, dc.pos); st4(c, r + 10, dc.rot);
  }
}

// ---- phase B2 (pbd): apply joint deltas (+ velocity projection in half A) ---
// integrators.py:103-110 then :122-146; system.py:277-281 / :296-297.
__device__ __forceinline__ void phase_body_joint_update(const Ctx& c, const BxHeader& h,
                                                     bool project) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  const int ppos = c.cur ? h.r_pos0 : h.r_pos1, prot = c.cur ? h.r_rot0 : h.r_rot1;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    V3 dpos = v3(0.f, 0.f, 0.f);
    Q4 drot = q4(0.f, 0.f, 0.f, 0.f);
    const int32_t* adj = c.ti + h.io_adj + bi[BI_JADJ_OFF];
    const int n = bi[BI_JADJ_N];
    for (int i = 0; i < n; i++) {
      int e = adj[i];
      int r = h.r_js + 14 * (e >> 1) + 7 * (e & 1);
      dpos = dpos + ld3(c, r);
      drot = drot + ld4(c, r + 3);
    }
    V3 rm = tf3(bf + BF_ROT_MASK);
    V3 pos = ld3(c, rpos + 3 * b) + dpos * tf3(bf + BF_POS_MASK);
    Q4 rot = ld4(c, rrot + 4 * b);
    rot = q4(rot.w + drot.w, rot.x + drot.x * rm.x, rot.y + drot.y * rm.y, rot.z + drot.z * rm.z);
    if (project) {
      V3 vel, ang;
      velocity_projection(h, bf, pos, ld3(c, ppos + 3 * b), ld4(c, prot + 4 * b), &rot, &vel, &ang);
      st3(c, h.r_vel + 3 * b, vel);
      st3(c, h.r_ang + 3 * b, ang);
    }
    st3(c, rpos + 3 * b, pos);
    st4(c, rrot + 4 * b, rot);
  }
}

// ---- contacts -----------------------------------------------------------------
struct Contact { V3 pos, vel, normal; float pen; };

// geometry.py:360-374
BXD V3 closest_segment_point_and_dist(V3 a, V3 b, V3 pt, float* d2) {
  V3 ab = b - a;
  float t = bx_div(dot(pt - a, ab), dot(ab, ab) + 1e-6f);
  V3 p = a + clipf(t, 0.f, 1.f) * ab;
  V3 v = pt - p;
  *d2 = dot(v, v);
  return p;
}

// geometry.py:394-459 closest_segment_to_segment_points
BXD void closest_seg_seg(V3 a0, V3 a1, V3 b0, V3 b1, V3* best_a, V3* best_b) {
  V3 dir_a = a1 - a0;
  float len_a = safe_norm(dir_a);
  len_a += 1e-6f * (len_a == 0.f ? 1.f : 0.f);
  dir_a = vdiv(dir_a, len_a);
  float half_a = len_a * 0.5f;
  V3 dir_b = b1 - b0;
  float len_b = safe_norm(dir_b);
  len_b += 1e-6f * (len_b == 0.f ? 1.f : 0.f);
  dir_b = vdiv(dir_b, len_b);
  float half_b = len_b * 0.5f;
  V3 a_mid = a0 + dir_a * half_a, b_mid = b0 + dir_b * half_b;
  V3 trans = a_mid - b_mid;
  float dab = dot(dir_a, dir_b), dat = dot(dir_a, trans), dbt = dot(dir_b, trans);
  float denom = 1.f - dab * dab;
  float ta0 = bx_div(-dat + dab * dbt, denom + 1e-6f);
  float tb0 = dbt + ta0 * dab;
  float ta = clipf(ta0, -half_a, half_a), tb = clipf(tb0, -half_b, half_b);
  V3 ba = a_mid + dir_a * ta, bb = b_mid + dir_b * tb;
  float d1, d2;
  V3 na = closest_segment_point_and_dist(a0, a1, bb, &d1);
  V3 nb = closest_segment_point_and_dist(b0, b1, ba, &d2);
  bool first = d1 < d2;
  *best_a = first ? na : ba;
  *best_b = first ? bb : nb;
}

BXD float w_of(float inv_mass, V3 inv_i, V3 r, V3 n) {
  V3 cr = cross(r, n);
  return inv_mass + dot(cr, inv_i * cr);
}

// contact function of one contact row (colliders.py:667-680 box_plane,
// :744-759 capsule_plane, :805-819 capsule_capsule).
BXD Contact contact_fn(const Ctx& c, const BxHeader& h, int kind, const float* cf,
                       V3 pa, Q4 ra, V3 va, V3 wa, V3 pb, Q4 rb, V3 vb, V3 wb) {
  Contact ct;
  if (kind == BX_PAIR_CAPSULE_CAPSULE) {
    V3 ca = pa + rotate(tf3(cf + CF_PTS + 3), ra), ea = rotate(tf3(cf + CF_PTS), ra);
    V3 cb = pb + rotate(tf3(cf + CF_PTS + 9), rb), eb = rotate(tf3(cf + CF_PTS + 6), rb);
    V3 best_a, best_b;
    closest_seg_seg(ca + ea, ca - ea, cb + eb, cb - eb, &best_a, &best_b);
    V3 pv = best_a - best_b;
    float dist = safe_norm(pv);
    ct.normal = vdiv(pv, 1e-6f + dist);
    ct.pen = cf[CF_RADIUS_A] + cf[CF_RADIUS_B] - dist;
    ct.pos = vdiv(best_a + best_b, 2.f);
    ct.vel = (va + cross(wa, ct.pos - pa)) - (vb + cross(wb, ct.pos - pb));
  } else {
    V3 n = rotate(v3(0.f, 0.f, 1.f), rb);
    V3 off = rotate(tf3(cf + CF_PTS), ra);
    if (kind == BX_PAIR_CAPSULE_PLANE) {
      ct.pos = (pa + off) - n * cf[CF_RADIUS_A];
      ct.vel = va + cross(wa, ct.pos - pa);
    } else {  // box corner: QP.to_world base.py:112-125
      ct.pos = pa + off;
      ct.vel = va + cross(wa, off);
    }
    ct.normal = n;
    ct.pen = dot(pb - ct.pos, n);
  }
  return ct;
}

// ---- phase C1 (pbd): contact_fn + _position_contact for every contact row ------
// colliders.py:198-240, OneWay :306-377, TwoWay :495-582.  Row k keeps its
// contact (pos3 normal3 pen dlambda) in rows CT+8k and its deltas (a: pos3 rot4
// [, b: pos3 rot4]) in rows JS+delta_off(k).
__device__ __forceinline__ void phase_contacts_position(const Ctx& c, const BxHeader& h) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  const int ppos = c.cur ? h.r_pos0 : h.r_pos1, prot = c.cur ? h.r_rot0 : h.r_rot1;
  for (int k = c.warp; k < h.nk; k += c.nwarp) {
    const int32_t* ci = c.ti + h.io_contact + k * CI_STRIDE;
    const float* cf = c.tf + h.fo_contact + k * CF_STRIDE;
    const int a = ci[CI_BODY_A], b = ci[CI_BODY_B];
    const bool two_way = ci[CI_TWO_WAY] != 0;
    const float* bfa = c.tf + h.fo_body + a * BF_STRIDE;
    const float* bfb = c.tf + h.fo_body + b * BF_STRIDE;
    V3 pa = ld3(c, rpos + 3 * a), pb = ld3(c, rpos + 3 * b);
    Q4 ra = ld4(c, rrot + 4 * a), rb = ld4(c, rrot + 4 * b);
    Contact ct = contact_fn(c, h, ci[CI_KIND], cf, pa, ra, ld3(c, h.r_vel + 3 * a),
                            ld3(c, h.r_ang + 3 * a), pb, rb, ld3(c, h.r_vel + 3 * b),
                            ld3(c, h.r_ang + 3 * b));
    V3 iia = tf3(bfa + BF_INV_INERTIA);
    const float ima = bfa[BF_INV_MASS], ma = bfa[BF_MASS];
    const float cs = h.collide_scale;
    V3 n = ct.normal;
    const int rd = h.r_js + ci[CI_DELTA_OFF];
    float dl_out;
    if (!two_way) {
      V3 pos_c = ct.pos + n * ct.pen;
      V3 dx = ct.pos - pos_c;
      V3 rp = ct.pos - pa;
      float cc = dot(dx, n);
      float w1 = w_of(ima, iia, rp, n);
      float dl = bx_div(-cc, w1 + 1e-6f);
      float cm = cc < 0.f ? 1.f : 0.f;
      V3 p = (dl * n) * cm;
      V3 dpos = cs * vdiv(p, ma);
      Q4 drot = cs * (.5f * vec_quat_mul(iia * cross(rp, p), ra));
      // static friction
      V3 r1 = rotate(rp, qconj(ra));
      V3 p1bar = ld3(c, ppos + 3 * a) + rotate(r1, ld4(c, prot + 4 * a));
      V3 deltap = ct.pos - p1bar;
      V3 dpt = deltap - dot(deltap, n) * n;
      float c2 = safe_norm(dpt);
      V3 n2 = vdiv(dpt, c2 + 1e-6f);
      float w1t = w_of(ima, iia, rp, n2);
      float dlt = bx_div(-c2, w1t + 0.f);
      float sm = fabsf(dlt) < fabsf(cf[CF_FRICTION] * dl) ? 1.f : 0.f;
      V3 pt = ((dlt * n2) * sm) * cm;
      dpos = dpos + cs * vdiv(pt, ma);
      drot = drot + cs * (.5f * vec_quat_mul(iia * cross(rp, pt), ra));
      st3(c, rd, dpos); st4(c, rd + 3, drot);
      dl_out = dl * cm;
    } else {
      V3 iib = tf3(bfb + BF_INV_INERTIA);
      const float imb = bfb[BF_INV_MASS], mb = bfb[BF_MASS];
      V3 half = vdiv(n * ct.pen, 2.f);
      V3 rp = (ct.pos - half) - pa, rc = (ct.pos + half) - pb;
      float cc = -ct.pen;
      float w1 = w_of(ima, iia, rp, n), w2 = w_of(imb, iib, rc, n);
      float dl = bx_div(-cc, w1 + w2 + 1e-6f);
      float cm = cc < 0.f ? 1.f : 0.f;
      V3 p = (dl * n) * cm;
      V3 dpa = cs * vdiv(p, ma), dpb = cs * vdiv(-p, mb);
      Q4 dra = cs * (.5f * vec_quat_mul(iia * cross(rp, p), ra));
      Q4 drb = cs * (-.5f * vec_quat_mul(iib * cross(rc, p), rb));
      // static friction
      V3 xa = ct.pos - pa, xb = ct.pos - pb;
      V3 r1 = rotate(xa, qconj(ra)), r2 = rotate(xb, qconj(rb));
      V3 p1bar = ld3(c, ppos + 3 * a) + rotate(r1, ld4(c, prot + 4 * a));
      V3 p2bar = ld3(c, ppos + 3 * b) + rotate(r2, ld4(c, prot + 4 * b));
      V3 deltap = (ct.pos - p1bar) - (ct.pos - p2bar);
      V3 dpt = deltap - dot(deltap, n) * n;
      float c2 = safe_norm(dpt);
      V3 n2 = vdiv(dpt, c2 + 1e-6f);
      float w1t = w_of(ima, iia, xa, n2), w2t = w_of(imb, iib, xb, n2);
      float dlt = bx_div(-c2, w1t + w2t);
      float sm = fabsf(dlt) < fabsf(dl) ? 1.f : 0.f;
      V3 pt = ((dlt * n2) * sm) * cm;
      dpa = dpa + cs * vdiv(pt, ma);
      dpb = dpb + cs * vdiv(-pt, mb);
      dra = dra + cs * (.5f * vec_quat_mul(iia * cross(xa, pt), ra));
      drb = drb + cs * (.5f * vec_quat_mul(iib * cross(xb, -pt), rb));
      st3(c, rd, dpa); st4(c, rd + 3, dra);
      st3(c, rd + 7, dpb); st4(c, rd + 10, drb);
      dl_out = dl;  // un-masked (colliders.py:580)
    }
    const int rc_ = h.r_ct + 8 * k;
    st3(c, rc_, ct.pos); st3(c, rc_ + 3, ct.normal);
    ROW(c, rc_ + 6) = ct.pen;
    ROW(c, rc_ + 7) = dl_out;
  }
}

// Per-body scatter-and-average of the contact rows' deltas, one collider group
// at a time (colliders.py:187-196, 231-240): sum rows, count rows with any
// non-zero linear part, divide by (eps + count), then sum over groups.
// W = 7 (Q deltas: pos3 rot4) or 6 (P deltas: vel3 ang3).
template <int W>
BXD void gather_contact_deltas(const Ctx& c, const BxHeader& h, const int32_t* bi,
                               float eps, float* out) {
#pragma unroll
  for (int q = 0; q < W; q++) out[q] = 0.f;
  const int32_t* adj = c.ti + h.io_adj + bi[BI_CADJ_OFF];
  const int n = bi[BI_CADJ_N];
  int i = 0;
  while (i < n) {
    const int g = (adj[i] >> 1) & 7;
    float s[W];
#pragma unroll
    for (int q = 0; q < W; q++) s[q] = 0.f;
    float cnt = 0.f;
    for (; i < n && ((adj[i] >> 1) & 7) == g; i++) {
      const int e = adj[i];
      const int32_t* ci = c.ti + h.io_contact + (e >> 4) * CI_STRIDE;
      const int r = h.r_js + ci[CI_DELTA_OFF] + (e & 1) * W;
      float v[W];
#pragma unroll
      for (int q = 0; q < W; q++) v[q] = ROW(c, r + q);
      cnt += (v[0] != 0.f || v[1] != 0.f || v[2] != 0.f) ? 1.f : 0.f;
#pragma unroll
      for (int q = 0; q < W; q++) s[q] += v[q];
    }
    const float d = eps + cnt, rd = bx_rcp(d);
#pragma unroll
    for (int q = 0; q < W; q++) out[q] += bx_mul_rcp(s[q], d, rd);
  }
}

// ---- phase B3 (pbd): apply contact deltas, keep qp_right_before's velocities,
// velocity projection.  system.py:299-307.
__device__ __forceinline__ void phase_body_contact_update(const Ctx& c, const BxHeader& h) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  const int ppos = c.cur ? h.r_pos0 : h.r_pos1, prot = c.cur ? h.r_rot0 : h.r_rot1;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    V3 pos = ld3(c, rpos + 3 * b);
    Q4 rot = ld4(c, rrot + 4 * b);
    if (bi[BI_CADJ_N] > 0) {
      float d[7];
      gather_contact_deltas<7>(c, h, bi, 1e-6f, d);
      V3 rm = tf3(bf + BF_ROT_MASK);
      pos = pos + v3(d[0], d[1], d[2]) * tf3(bf + BF_POS_MASK);
      rot = q4(rot.w + d[3], rot.x + d[4] * rm.x, rot.y + d[5] * rm.y, rot.z + d[6] * rm.z);
      st3(c, rpos + 3 * b, pos);
    }
    if (bi[BI_CSLOT] >= 0) {  // qp_right_before.vel / .ang (system.py:306)
      const int r = h.r_pre + 6 * bi[BI_CSLOT];
      st3(c, r, ld3(c, h.r_vel + 3 * b));
      st3(c, r + 3, ld3(c, h.r_ang + 3 * b));
    }
    V3 vel, ang;
    velocity_projection(h, bf, pos, ld3(c, ppos + 3 * b), ld4(c, prot + 4 * b), &rot, &vel, &ang);
    st3(c, h.r_vel + 3 * b, vel);
    st3(c, h.r_ang + 3 * b, ang);
    st4(c, rrot + 4 * b, rot);
  }
}

// ---- phase C2 (pbd): _velocity_contact for every contact row ------------------
// colliders.py:155-196, OneWay :379-442, TwoWay :584-658.  Deltas (a: vel3 ang3
// [, b: vel3 ang3]) go to rows JS+delta_off(k).
__device__ __forceinline__ void phase_contacts_velocity(const Ctx& c, const BxHeader& h) {
  const int rpos = c.cur ? h.r_pos1 : h.r_pos0;
  for (int k = c.warp; k < h.nk; k += c.nwarp) {
    const int32_t* ci = c.ti + h.io_contact + k * CI_STRIDE;
    const float* cf = c.tf + h.fo_contact + k * CF_STRIDE;
    const int a = ci[CI_BODY_A], b = ci[CI_BODY_B];
    const bool two_way = ci[CI_TWO_WAY] != 0;
    const int32_t* bia = c.ti + h.io_body + a * BI_STRIDE;
    const float* bfa = c.tf + h.fo_body + a * BF_STRIDE;
    const int rc_ = h.r_ct + 8 * k;
    V3 cpos = ld3(c, rc_), n = ld3(c, rc_ + 3);
    const float pen = ROW(c, rc_ + 6), dlambda = ROW(c, rc_ + 7);
    V3 pa = ld3(c, rpos + 3 * a);
    V3 ra = cpos - pa;
    V3 iia = tf3(bfa + BF_INV_INERTIA);
    const float ima = bfa[BF_INV_MASS], ma = bfa[BF_MASS];
    const int rpre_a = h.r_pre + 6 * bia[BI_CSLOT];
    const int rd = h.r_js + ci[CI_DELTA_OFF];
    const float stat = pen > 0.f ? 1.f : 0.f;
    if (!two_way) {
      V3 rel = ld3(c, h.r_vel + 3 * a) + cross(ld3(c, h.r_ang + 3 * a), ra);
      float v_n = dot(rel, n);
      V3 v_t = rel - n * v_n;
      float vtn = safe_norm(v_t);
      V3 vdir = vdiv(v_t, 1e-6f + vtn);
      float lim = (cf[CF_FRICTION] * fabsf(dlambda)) * h.inv_2h;
      V3 dvel = (-vdir) * fminf(lim, vtn);
      V3 angw = cross(ra, vdir);
      float w = ima + dot(angw, angw);  // no inertia here (colliders.py:405-406)
      V3 p_dyn = vdiv(dvel, w + 1e-6f);
      // restitution; qp_right_before has the same pos as now
      V3 rel_old = ld3(c, rpre_a) + cross(ld3(c, rpre_a + 3), ra);
      float v_n_old = dot(rel_old, n);
      V3 dx = n * (-v_n - fminf(cf[CF_ELASTICITY] * v_n_old, 0.f));
      float cc = safe_norm(dx);
      V3 n2 = vdiv(dx, cc + 1e-6f);
      float w1 = w_of(ima, iia, ra, n2);
      float dlr = bx_div(cc, w1 + 1e-6f);
      float sink = v_n_old <= -h.vth ? 1.f : 0.f;
      V3 p = (((dlr * n2) * sink) + p_dyn) * stat;
      st3(c, rd, vdiv(p, ma));
      st3(c, rd + 3, cross(iia * ra, p));
    } else {
      const int32_t* bib = c.ti + h.io_body + b * BI_STRIDE;
      const float* bfb = c.tf + h.fo_body + b * BF_STRIDE;
      V3 iib = tf3(bfb + BF_INV_INERTIA);
      const float imb = bfb[BF_INV_MASS], mb = bfb[BF_MASS];
      const int rpre_b = h.r_pre + 6 * bib[BI_CSLOT];
      V3 pb = ld3(c, rpos + 3 * b);
      V3 rb = cpos - pb;
      V3 rel = (ld3(c, h.r_vel + 3 * a) + cross(ld3(c, h.r_ang + 3 * a), ra)) -
               (ld3(c, h.r_vel + 3 * b) + cross(ld3(c, h.r_ang + 3 * b), rb));
      float v_n = dot(rel, n);
      V3 v_t = rel - n * v_n;
      float vtn = safe_norm(v_t);
      V3 vdir = vdiv(v_t, 1e-6f + vtn);
      float lim = (cf[CF_FRICTION] * fabsf(dlambda)) * h.inv_2h;
      V3 dvel = (-vdir) * fminf(lim, vtn);
      float w1 = w_of(ima, iia, ra, vdir), w2 = w_of(imb, iib, rb, vdir);
      V3 p_dyn = vdiv(dvel, w1 + w2 + 1e-6f);
      V3 rel_old = (ld3(c, rpre_a) + cross(ld3(c, rpre_a + 3), ra)) -
                   (ld3(c, rpre_b) + cross(ld3(c, rpre_b + 3), rb));
      float v_n_old = dot(rel_old, n);
      V3 dx = n * (-v_n - fminf(cf[CF_ELASTICITY] * v_n_old, 0.f));
      V3 rc2 = (cpos + n * pen) - pb;
      float cc = safe_norm(dx);
      V3 n2 = vdiv(dx, cc + 1e-6f);
      float w1r = w_of(ima, iia, ra, n2), w2r = w_of(imb, iib, rc2, n2);
      float dlr = bx_div(cc, w1r + w2r + 1e-6f);
      float sink = v_n_old <= 0.f ? 1.f : 0.f;
      V3 p = (((dlr * n2) * sink) + p_dyn) * stat;
      st3(c, rd, vdiv(p, ma));
      st3(c, rd + 3, cross(iia * ra, p));
      st3(c, rd + 6, vdiv(-p, mb));
      st3(c, rd + 9, cross(iib * rb, -p));
    }
  }
}

// ---- phase B4: velocity-level contact update + Info.contact -----------------
// integrators.py:97-101; system.py:313,315 (pbd) / :365,367 (legacy_spring).
__device__ __forceinline__ void phase_body_velocity_update(const Ctx& c, const BxHeader& h,
                                                        float eps) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    if (bi[BI_CADJ_N] == 0) continue;
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    float d[6];
    gather_contact_deltas<6>(c, h, bi, eps, d);
    V3 dv = v3(d[0], d[1], d[2]), da = v3(d[3], d[4], d[5]);
    st3(c, h.r_vel + 3 * b, (ld3(c, h.r_vel + 3 * b) + dv) * tf3(bf + BF_POS_MASK));
    st3(c, h.r_ang + 3 * b, (ld3(c, h.r_ang + 3 * b) + da) * tf3(bf + BF_ROT_MASK));
    const int r = h.r_ic + 6 * bi[BI_CSLOT];
    st3(c, r, ld3(c, r) + dv);
    st3(c, r + 3, ld3(c, r + 3) + da);
  }
}

// ======================= legacy_spring (system.py:342-377) =====================
// Body.impulse bodies.py:46-59
BXD void body_impulse(float mass, V3 inv_i, V3 bpos, V3 imp, V3 at, V3* dvel, V3* dang) {
  *dvel = vdiv(imp, mass);
  *dang = inv_i * cross(at - bpos, imp);
}

// phase S1: kinetic step for every body, in place (system.py:352).
__device__ __forceinline__ void phase_body_kinetic(const Ctx& c, const BxHeader& h) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    V3 pos = ld3(c, h.r_pos0 + 3 * b);
    Q4 rot = ld4(c, h.r_rot0 + 4 * b);
    kinetic(h, bf, ld3(c, h.r_vel + 3 * b), ld3(c, h.r_ang + 3 * b), &pos, &rot);
    st3(c, h.r_pos0 + 3 * b, pos);
    st4(c, h.r_rot0 + 4 * b, rot);
  }
}

// phase S2: spring joints (spring_joints.py:89-303) -> slot rows
// JS+3na+12j: parent vel3 ang3, child vel3 ang3.
__device__ __forceinline__ void phase_joints_spring(const Ctx& c, const BxHeader& h) {
  for (int j = c.warp; j < h.nj; j += c.nwarp) {
    const int32_t* ji = c.ti + h.io_joint + j * JI_STRIDE;
    const float* jf = c.tf + h.fo_joint + j * JF_STRIDE;
    const int kind = ji[JI_KIND], bp = ji[JI_BODY_P], bc = ji[JI_BODY_C];
    const float* bfp = c.tf + h.fo_body + bp * BF_STRIDE;
    const float* bfc = c.tf + h.fo_body + bc * BF_STRIDE;
    V3 pos_p = ld3(c, h.r_pos0 + 3 * bp), pos_c = ld3(c, h.r_pos0 + 3 * bc);
    Q4 rp = ld4(c, h.r_rot0 + 4 * bp), rc = ld4(c, h.r_rot0 + 4 * bc);
    V3 ang_p = ld3(c, h.r_ang + 3 * bp), ang_c = ld3(c, h.r_ang + 3 * bc);
    V3 ii_p = tf3(bfp + BF_INV_INERTIA), ii_c = tf3(bfc + BF_INV_INERTIA);
    // QP.to_world base.py:112-125
    V3 op = rotate(tf3(jf + JF_OFF_P), rp), oc = rotate(tf3(jf + JF_OFF_C), rc);
    V3 wp = pos_p + op, wc = pos_c + oc;
    V3 vp = ld3(c, h.r_vel + 3 * bp) + cross(ang_p, op);
    V3 vc = ld3(c, h.r_vel + 3 * bc) + cross(ang_c, oc);
    V3 imp = (wp - wc) * jf[JF_STIFFNESS] + jf[JF_SPRING_DAMP] * (vp - vc);
    V3 dvp, dap, dvc, dac;
    body_impulse(bfp[BF_MASS], ii_p, pos_p, -imp, wp, &dvp, &dap);
    body_impulse(bfc[BF_MASS], ii_c, pos_c, imp, wc, &dvc, &dac);
    V3 axis[3];
    float angle[3], dang[3];
    joint_axis_angle(kind, jf, rp, rc, axis, angle);
    const int dof = joint_dof(kind);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      float lo = jf[JF_LIMIT + 2 * d], hi = jf[JF_LIMIT + 2 * d + 1];
      float v = 0.f;
      if (d < dof) {
        v = angle[d] < lo ? lo - angle[d] : 0.f;
        v = angle[d] > hi ? hi - angle[d] : v;
      }
      dang[d] = v;
    }
    const float ls = jf[JF_LIMIT_STRENGTH];
    V3 tq;
    if (kind == BX_JOINT_SPRING_REVOLUTE) {  // spring_joints.py:127-148
      V3 axis_c = rotate(tf3(jf + JF_AXIS_C), rc);
      tq = jf[JF_STIFFNESS] * cross(axis[0], axis_c);
      tq = tq - (ls * axis[0]) * dang[0];
    } else if (kind == BX_JOINT_SPRING_UNIVERSAL) {  // :178-210
      V3 proj = axis[1] - dot(axis[1], axis[0]) * axis[0];
      proj = vdiv(proj, safe_norm(proj));
      tq = (ls / 5.f) * cross(proj, axis[1]);
      tq = tq - ls * (axis[0] * dang[0] + axis[1] * dang[1]);
    } else {  // spring Spherical :249-274
      tq = (-ls) * ((axis[0] * dang[0] + axis[1] * dang[1]) + axis[2] * dang[2]);
    }
    tq = tq - jf[JF_ANG_DAMP] * (ang_p - ang_c);
    dap = dap + ii_p * tq;
    dac = dac + (-ii_c) * tq;
    const int r = h.r_js + 3 * h.na + 12 * j;
    st3(c, r, dvp); st3(c, r + 3, dap); st3(c, r + 6, dvc); st3(c, r + 9, dac);
  }
}

// phase S3: acceleration update from joints + actuators + forces; Info.joint and
// Info.actuator accumulation (system.py:356-359,367).
__device__ __forceinline__ void phase_body_acc_spring(const Ctx& c, const BxHeader& h) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float* bf = c.tf + h.fo_body + b * BF_STRIDE;
    V3 dpa, unused, fv, fa;
    body_accels(c, h, b, bi, bf, false, &dpa, &unused, &fv, &fa);
    V3 jv = v3(0.f, 0.f, 0.f), ja = v3(0.f, 0.f, 0.f);
    const int32_t* adj = c.ti + h.io_adj + bi[BI_JADJ_OFF];
    const int n = bi[BI_JADJ_N];
    for (int i = 0; i < n; i++) {
      int e = adj[i];
      int r = h.r_js + 3 * h.na + 12 * (e >> 1) + 6 * (e & 1);
      jv = jv + ld3(c, r);
      ja = ja + ld3(c, r + 3);
    }
    V3 g = v3(h.gravity[0], h.gravity[1], h.gravity[2]);
    V3 acc_v = jv + fv, acc_a = (ja + dpa) + fa;
    V3 vel = (h.vel_damp * ld3(c, h.r_vel + 3 * b) + (acc_v + g) * h.dt) * tf3(bf + BF_POS_MASK);
    V3 ang = (h.ang_damp * ld3(c, h.r_ang + 3 * b) + acc_a * h.dt) * tf3(bf + BF_ROT_MASK);
    st3(c, h.r_vel + 3 * b, vel);
    st3(c, h.r_ang + 3 * b, ang);
    const int rj = h.r_ij + 6 * b;
    st3(c, rj, ld3(c, rj) + jv);
    st3(c, rj + 3, ld3(c, rj + 3) + ja);
    if (bi[BI_ASLOT] >= 0) {
      int r = h.r_ia + 3 * bi[BI_ASLOT];
      st3(c, r, ld3(c, r) + dpa);
    }
  }
}

// phase S4: contact_fn + velocity-level _contact (colliders.py:116-153,
// OneWay :267-304, TwoWay :449-493).
__device__ __forceinline__ void phase_contacts_spring(const Ctx& c, const BxHeader& h) {
  for (int k = c.warp; k < h.nk; k += c.nwarp) {
    const int32_t* ci = c.ti + h.io_contact + k * CI_STRIDE;
    const float* cf = c.tf + h.fo_contact + k * CF_STRIDE;
    const int a = ci[CI_BODY_A], b = ci[CI_BODY_B];
    const bool two_way = ci[CI_TWO_WAY] != 0;
    const float* bfa = c.tf + h.fo_body + a * BF_STRIDE;
    const float* bfb = c.tf + h.fo_body + b * BF_STRIDE;
    V3 pa = ld3(c, h.r_pos0 + 3 * a), pb = ld3(c, h.r_pos0 + 3 * b);
    Contact ct = contact_fn(c, h, ci[CI_KIND], cf, pa, ld4(c, h.r_rot0 + 4 * a),
                            ld3(c, h.r_vel + 3 * a), ld3(c, h.r_ang + 3 * a), pb,
                            ld4(c, h.r_rot0 + 4 * b), ld3(c, h.r_vel + 3 * b),
                            ld3(c, h.r_ang + 3 * b));
    V3 iia = tf3(bfa + BF_INV_INERTIA), iib = tf3(bfb + BF_INV_INERTIA);
    V3 n = ct.normal;
    V3 ra = ct.pos - pa, rb = ct.pos - pb;
    float baum = h.erp * ct.pen;
    float nv = dot(n, ct.vel);
    V3 c1 = cross(iia * cross(ra, n), ra);
    float ang, denom;
    if (two_way) {
      V3 c2 = cross(iib * cross(rb, n), rb);
      ang = dot(n, c1 + c2);
      denom = bx_div(1.f, bfa[BF_MASS]) + bx_div(1.f, bfb[BF_MASS]) + ang;
    } else {
      ang = dot(n, c1);
      denom = bx_div(1.f, bfa[BF_MASS]) + ang;
    }
    float imp = bx_div(-1.f * (1.f + cf[CF_ELASTICITY]) * nv + baum, denom);
    V3 vd = ct.vel - nv * n;
    float nvd = safe_norm(vd);
    float impd = fminf(bx_div(nvd, denom), cf[CF_FRICTION] * imp);
    V3 dir = vdiv(vd, 1e-6f + nvd);
    float apn = (ct.pen > 0.f && nv < 0.f && imp > 0.f) ? 1.f : 0.f;
    float apd = apn * (nvd > 0.01f ? 1.f : 0.f);
    V3 dvn, dan, dvd, dad;
    body_impulse(bfa[BF_MASS], iia, pa, imp * n, ct.pos, &dvn, &dan);
    body_impulse(bfa[BF_MASS], iia, pa, (-impd) * dir, ct.pos, &dvd, &dad);
    const int rd = h.r_js + ci[CI_DELTA_OFF];
    st3(c, rd, dvn * apn + dvd * apd);
    st3(c, rd + 3, dan * apn + dad * apd);
    if (two_way) {
      body_impulse(bfb[BF_MASS], iib, pb, (-imp) * n, ct.pos, &dvn, &dan);
      body_impulse(bfb[BF_MASS], iib, pb, impd * dir, ct.pos, &dvd, &dad);
      st3(c, rd + 6, dvn * apn + dvd * apd);
      st3(c, rd + 9, dan * apn + dad * apd);
    }
    const int rc_ = h.r_ct + 8 * k;
    st3(c, rc_, ct.pos); st3(c, rc_ + 3, ct.normal);
    ROW(c, rc_ + 6) = ct.pen;
  }
}

// ============================== the kernel ====================================
extern __shared__ __align__(16) float bx_smem[];

__device__ __forceinline__ void zero_rows(float* sm, int row0, int nrows) {
  for (int i = threadIdx.x; i < nrows * RS; i += blockDim.x) sm[row0 * RS + i] = 0.f;
}

// blockDim.x = 32 * W (W warps share one tile of 32 environments).
template <int MODE>
__global__ void __launch_bounds__(512, 1)
bx_step_kernel(const __grid_constant__ BxKernelParams p) {
  const BxHeader& h = p.h;
  float* sm = bx_smem;
  float* tf = sm + h.n_rows * RS;
  int32_t* ti = reinterpret_cast<int32_t*>(tf + h.n_floats);
  const int64_t env0 = static_cast<int64_t>(blockIdx.x) * BX_ENVS_PER_CTA;
  const int nenv = static_cast<int>((p.num_envs - env0) < BX_ENVS_PER_CTA ? (p.num_envs - env0) : BX_ENVS_PER_CTA);

  for (int i = threadIdx.x; i < h.n_floats; i += blockDim.x) tf[i] = __ldg(p.tab_f + i);
  for (int i = threadIdx.x; i < h.n_ints; i += blockDim.x) ti[i] = __ldg(p.tab_i + i);
  if (nenv < BX_ENVS_PER_CTA) {
    // partial tile: give the idle env slots a benign state (identity rotations)
    for (int i = threadIdx.x; i < h.n_rows * RS; i += blockDim.x) sm[i] = 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < h.nb * BX_ENVS_PER_CTA; i += blockDim.x) {
      int b = i / BX_ENVS_PER_CTA, e = i - b * BX_ENVS_PER_CTA;
      sm[(h.r_rot0 + 4 * b) * RS + e] = 1.f;
    }
    __syncthreads();
  }
  load_rows(sm, p.io.pos, env0, nenv, 3 * h.nb, h.r_pos0);
  load_rows(sm, p.io.rot, env0, nenv, 4 * h.nb, h.r_rot0);
  load_rows(sm, p.io.vel, env0, nenv, 3 * h.nb, h.r_vel);
  load_rows(sm, p.io.ang, env0, nenv, 3 * h.nb, h.r_ang);

  Ctx c;
  c.lane = threadIdx.x & 31;
  c.warp = threadIdx.x >> 5;
  c.nwarp = blockDim.x >> 5;
  c.st = sm + c.lane;
  c.tf = tf;
  c.ti = ti;
  c.cur = 0;

  for (int step = 0; step < p.num_steps; step++) {
    __syncthreads();
    load_rows(sm, p.io.act + static_cast<int64_t>(step) * p.num_envs * h.act_size, env0,
              nenv, h.act_size, h.r_act);
    // zero_info (system.py:76-84): Info accumulators start at 0 every step
    zero_rows(sm, h.r_ic, 6 * h.ncb + 3 * h.nab + (MODE == BX_MODE_LEGACY_SPRING ? 6 * h.nb : 0));
    __syncthreads();
    if constexpr (MODE == BX_MODE_PBD) {
      for (int it = 0; it < h.iters; it++) {
#pragma unroll 1
        for (int half = 0; half < 2; half++) {
          phase_torques(c, h, true);
          __syncthreads();
          phase_body_acc_kinetic(c, h, half == 1);
          c.cur ^= 1;
          __syncthreads();
          phase_joints_pbd(c, h);
          __syncthreads();
          phase_body_joint_update(c, h, half == 0);
          __syncthreads();
          if (half == 1) {
            phase_contacts_position(c, h);
            __syncthreads();
            phase_body_contact_update(c, h);
            __syncthreads();
            phase_contacts_velocity(c, h);
            __syncthreads();
            phase_body_velocity_update(c, h, 1e-6f);
            __syncthreads();
          }
        }
      }
    } else {
      for (int it = 0; it < h.iters; it++) {
        phase_body_kinetic(c, h);
        __syncthreads();
        phase_torques(c, h, false);
        phase_joints_spring(c, h);
        __syncthreads();
        phase_body_acc_spring(c, h);
        __syncthreads();
        phase_contacts_spring(c, h);
        __syncthreads();
        phase_body_velocity_update(c, h, 1e-8f);
        // same warp<->body mapping as phase_body_kinetic: no barrier needed
      }
    }
  }
  __syncthreads();

  const int rpos = c.cur ? h.r_pos1 : h.r_pos0, rrot = c.cur ? h.r_rot1 : h.r_rot0;
  store_rows(sm, p.io.pos_o, env0, nenv, 3 * h.nb, rpos, nullptr);
  store_rows(sm, p.io.rot_o, env0, nenv, 4 * h.nb, rrot, nullptr);
  store_rows(sm, p.io.vel_o, env0, nenv, 3 * h.nb, h.r_vel, nullptr);
  store_rows(sm, p.io.ang_o, env0, nenv, 3 * h.nb, h.r_ang, nullptr);
  store_rows(sm, p.io.contact_vel, env0, nenv, 3 * h.nb, 0, ti + h.m_cv);
  store_rows(sm, p.io.contact_ang, env0, nenv, 3 * h.nb, 0, ti + h.m_ca);
  store_rows(sm, p.io.actuator_ang, env0, nenv, 3 * h.nb, 0, ti + h.m_aa);
  store_const(p.io.actuator_vel, env0, nenv, 3 * h.nb, 0.f);
  if constexpr (MODE == BX_MODE_LEGACY_SPRING) {
    store_rows(sm, p.io.joint_vel, env0, nenv, 3 * h.nb, 0, ti + h.m_jv);
    store_rows(sm, p.io.joint_ang, env0, nenv, 3 * h.nb, 0, ti + h.m_ja);
  } else {
    store_const(p.io.joint_vel, env0, nenv, 3 * h.nb, 0.f);
    store_const(p.io.joint_ang, env0, nenv, 3 * h.nb, 0.f);
  }
  if (h.nk > 0) {
    if (h.iters > 0) {
      store_rows(sm, p.io.contact_pos, env0, nenv, 3 * h.nk, 0, ti + h.m_cp);
      store_rows(sm, p.io.contact_normal, env0, nenv, 3 * h.nk, 0, ti + h.m_cn);
      store_rows(sm, p.io.contact_penetration, env0, nenv, h.nk, 0, ti + h.m_pen);
    } else {  // zero_info: penetration = -1 (system.py:84)
      store_const(p.io.contact_pos, env0, nenv, 3 * h.nk, 0.f);
      store_const(p.io.contact_normal, env0, nenv, 3 * h.nk, 0.f);
      store_const(p.io.contact_penetration, env0, nenv, h.nk, -1.f);
    }
  }
}
