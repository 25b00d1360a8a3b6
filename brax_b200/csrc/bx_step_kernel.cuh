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
//     Info accumulators live in shared memory for the entire substep loop as
//     float4 "quads" laid out [quad row][env] (row stride 33 quads): a 3-vector
//     or a quaternion of one env is ONE 128-bit LDS/STS, conflict-free across
//     the 32 lanes, and the transposing global I/O is conflict-free too;
//   * the static tables are float4 / int4 records read with 128-bit broadcasts;
//   * HBM is touched once on the way in and once on the way out, with
//     coalesced 128-bit loads/stores of the boundary layout [B,N,3|4].
// FP32 CUDA-core bound by design; tensor cores do not apply (no contraction).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/brax_b200.h"
#include "bx_math.cuh"
#include "bx_params.h"

// Phase timing (tools/phase_timing.py builds with -DBX_PHASE_TIMING): thread 0
// of CTA 0 accumulates the cycles between consecutive barriers per phase id.
#ifdef BX_PHASE_TIMING
// dbg[0..15]: phase cycles of CTA 0 (barrier to barrier); dbg[16..31]: of those, the
// cycles warp 0 was busy before it arrived at the barrier; dbg[32 + 4 * cta ..]:
// {smid, start ns, end ns, cycles}
#define PT_INIT()                                                                  \
  long long pt_last = clock64();                                                   \
  const long long pt_c0 = pt_last;                                                 \
  long long pt_g0;                                                                 \
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(pt_g0))
#define PT(i)                                                        \
  do {                                                               \
    if (p.dbg && blockIdx.x == 0 && threadIdx.x == 0) {              \
      long long t_ = clock64();                                      \
      atomicAdd(reinterpret_cast<unsigned long long*>(p.dbg + (i)), static_cast<unsigned long long>(t_ - pt_last)); \
      pt_last = t_;                                                  \
    }                                                                \
  } while (0)
#define PTA(i)                                                       \
  do {                                                               \
    if (p.dbg && blockIdx.x == 0 && threadIdx.x == 0)                \
      atomicAdd(reinterpret_cast<unsigned long long*>(p.dbg + 16 + (i)), static_cast<unsigned long long>(clock64() - pt_last)); \
  } while (0)
#define BX_PTM_BASE (32 + 4 * 4096)
#define PTM0(c) do { if ((c).dbg && threadIdx.x == 0 && blockIdx.x == 0) (c).mark = clock64(); } while (0)
#define PTM(c, slot)                                                  \
  do {                                                                \
    if ((c).dbg && threadIdx.x == 0 && blockIdx.x == 0) {             \
      long long t_ = clock64();                                       \
      atomicAdd(reinterpret_cast<unsigned long long*>((c).dbg + BX_PTM_BASE + (c).moff + (slot)), static_cast<unsigned long long>(t_ - (c).mark)); \
      (c).mark = t_;                                                  \
    }                                                                 \
  } while (0)
#define PT_DONE()                                                                  \
  do {                                                                             \
    if (p.dbg && threadIdx.x == 0 && blockIdx.x < 4096) {                          \
      long long g1_;                                                               \
      unsigned sm_;                                                                \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1_));                      \
      asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_));                             \
      long long* r_ = p.dbg + 32 + 4 * blockIdx.x;                                 \
      r_[0] = sm_; r_[1] = pt_g0; r_[2] = g1_; r_[3] = clock64() - pt_c0;          \
    }                                                                              \
  } while (0)
#else
#define PT_INIT()
#define PT(i)
#define PTA(i)
#define PTM0(c)
#define PTM(c, slot)
#define PT_DONE()
#endif

#define QS BX_QSTRIDE

// Inside Kern<FEAT> the mask is the template parameter.
#define BX_FEAT FEAT
// "is x the case?" with the answer folded at compile time when the mask decides it
// the env epilogue phases (once per step, after the substep loop): inlined (shared-memory loads instead of the
// generic ones a separate function gets) unless -DBX_ENV_NOINLINE
#ifdef BX_ENV_NOINLINE
#define BX_ENV_INLINE __noinline__
#else
#define BX_ENV_INLINE __forceinline__
#endif
#define BXF_IS(cond, yes_mask, no_mask) \
  ((BX_FEAT & (no_mask)) == 0 ? true : ((BX_FEAT & (yes_mask)) == 0 ? false : (cond)))

struct Ctx {
  float4* q;          // shared state base + lane
  const float4* tf;   // float4 tables (shared)
  const int32_t* ti;  // int tables (shared)
  int warp, nwarp;
  int cur;            // ping-pong index of the current pos/rot quads (pbd)
  int damp_j0, damp_step;  // this warp's joint-damping items (phase T), set once per launch
#ifdef BX_PHASE_TIMING
  long long* dbg;     // within-item markers (PTM): dbg[BX_PTM_BASE + slot]
  mutable long long mark;
  int moff;           // marker slot offset (experiments)
#endif
};

// body block layout (quads): pbd {pos0 rot0 pos1 rot1 vel ang}, spring {pos rot vel ang}
template <int MODE> struct Lay {
  static constexpr int BQ = MODE == BX_MODE_PBD ? BX_BODY_QUADS_PBD : BX_BODY_QUADS_SPRING;
  static constexpr int VEL = BQ - 2, ANG = BQ - 1;
};

#define QROW(c, r) ((c).q[(r) * QS])

extern __shared__ __align__(16) float4 bx_smem[];

// All device code of the step lives in Kern<FEAT>: FEAT (a BXF_* mask) folds the
// feature tests at compile time.
#define BXS static __device__ __forceinline__
template <int FEAT>
struct Kern {
BXS V3 f3(float4 t) { return v3(t.x, t.y, t.z); }
BXS V3 ldv(const Ctx& c, int r) { return f3(QROW(c, r)); }
BXS Q4 ldq(const Ctx& c, int r) { float4 t = QROW(c, r); return q4(t.x, t.y, t.z, t.w); }
BXS void stv(const Ctx& c, int r, V3 v) { QROW(c, r) = make_float4(v.x, v.y, v.z, 0.f); }
BXS void stv4(const Ctx& c, int r, V3 v, float w) { QROW(c, r) = make_float4(v.x, v.y, v.z, w); }
BXS void stq(const Ctx& c, int r, Q4 q) { QROW(c, r) = make_float4(q.w, q.x, q.y, q.z); }
BXS float act_at(const Ctx& c, const BxHeader& h, int i) {
  return reinterpret_cast<const float*>(&QROW(c, h.q_act + (i >> 2)))[i & 3];
}
BXS int4 ldi4(const int32_t* p) { return *reinterpret_cast<const int4*>(p); }

// ---- global <-> shared transposing I/O ---------------------------------------
// g: [B, nper] row-major; this CTA owns envs [env0, env0+nenv).  Element (e, r)
// lives at shared float index map[r] + 4*e (map[r] < 0: constant `fill`).
BXS void seg_locate(const BxSeg& sg, int idx, int* e, int* r) {
  *e = sg.magic ? static_cast<int>(__umulhi(static_cast<unsigned>(idx), sg.magic)) : idx;
  *r = idx - *e * sg.nper;
}

// scalar fallback (ragged tiles / unaligned pointers)
template <bool LOAD>
static __device__ __noinline__ void io_segment_scalar(float* smf, const int32_t* ti, BxSeg sg, float* g,
                                               int64_t env0, int nenv) {
  float* base = g + env0 * sg.nper;
  const int32_t* map = ti + sg.map;
  const int total = nenv * sg.nper;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    int e, r;
    seg_locate(sg, idx, &e, &r);
    const int m = sg.map >= 0 ? map[r] : -1;
    if (LOAD) { if (m >= 0) smf[m + 4 * e] = base[idx]; }  // plain load: with PDL the producer may still run at CTA start
    else base[idx] = m >= 0 ? smf[m + 4 * e] : sg.fill;
  }
}

// ---- bulk asynchronous copies (TMA engine, no tensor map needed: every span of
// a 32-env tile is contiguous in HBM) ------------------------------------------
BXS uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
BXS void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
BXS void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
BXS void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "BX_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra BX_DONE;\n\t"
      "bra BX_WAIT;\n\t"
      "BX_DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
BXS void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
BXS void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
BXS void bulk_commit_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
BXS void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// Programmatic dependent launch (no-ops for a normally serialized launch): the next kernel of
// the stream may start launching once every CTA of this grid has passed pdl_trigger; a CTA of
// that kernel may touch what this grid wrote only after pdl_wait (= this grid complete, flushed).
BXS void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
BXS void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

BXS bool span_bulk_ok(const BxSeg& sg, const float* g, int64_t env0) {
  return (reinterpret_cast<uintptr_t>(g + env0 * sg.nper) & 15) == 0;
}

// One WARP moves one whole segment of a full tile, four consecutive elements per
// lane and iteration (128-bit on the global / staging side).  All warps run this
// same small loop on different segments at once.
// staging (global layout [env][nper]) -> state quads
BXS void unstage_warp(float* smf, const int32_t* ti, const BxSeg& sg, const float* stage, int lane) {
  const int32_t* map = ti + sg.map;
  const int n4 = BX_ENVS_PER_CTA / 4 * sg.nper;
  const float4* st4 = reinterpret_cast<const float4*>(stage);
#pragma unroll 1
  for (int i = lane; i < n4; i += 32) {
    int e, r;
    seg_locate(sg, i << 2, &e, &r);
    const float4 v = st4[i];
    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int m = map[r];
      if (m >= 0) smf[m + 4 * e] = vv[k];
      if (++r == sg.nper) { r = 0; e++; }
    }
  }
}
// state quads -> global (coalesced 128-bit stores); constant segments are a fill
// (lane, stride) = (lane, 32): one warp moves the segment; (threadIdx.x, blockDim.x): the
// whole CTA shares it (long segments: the env epilogue's obs)
BXS void store_warp(const float* smf, const int32_t* ti, const BxSeg& sg, float* gspan, int lane,
                    int stride = 32) {
  const int n4 = BX_ENVS_PER_CTA / 4 * sg.nper;
  float4* g4 = reinterpret_cast<float4*>(gspan);
  if (sg.map < 0) {
    const float4 f = make_float4(sg.fill, sg.fill, sg.fill, sg.fill);
    for (int i = lane; i < n4; i += stride) g4[i] = f;
    return;
  }
  const int32_t* map = ti + sg.map;
#pragma unroll 1
  for (int i = lane; i < n4; i += stride) {
    int e, r;
    seg_locate(sg, i << 2, &e, &r);
    float vv[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int m = map[r];
      vv[k] = m >= 0 ? smf[m + 4 * e] : sg.fill;
      if (++r == sg.nper) { r = 0; e++; }
    }
    g4[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
  }
}

// Joint float record stride and the limit directions {cos lo, sin lo, cos hi, sin hi} of dof d
// (bx_tables.h): systems with revolute joints only (the BXF_JC0 variants) keep the 10-quad record --
// their dof-0 directions take the place of the unused {lo2, hi2} quad -- so their tiles do not grow
// (ant stays under the three-tiles-per-SM limit); all others carry three more quads.
static constexpr int JQS = (BX_FEAT & (BXF_SPHERICAL | BXF_UNIVERSAL)) ? JQ_STRIDE : JQ_STRIDE_REVOLUTE;
BXS float4 limcs(const float4* jq, int d) {
  return (BX_FEAT & (BXF_SPHERICAL | BXF_UNIVERSAL)) ? jq[JQ_LIMCS + d] : jq[JQ_LIMIT + 1];
}

// ---- joints: axis / angle -------------------------------------------------
// Revolute.axis_angle joints.py:311-319 (spring_joints.py:150-157); mp / mc = rotmat(rot_p / rot_c)
BXS void axis_angle_revolute(const float4* jq, const M3& mp, const M3& mc, V3* axis, float* angle) {
  V3 ax = mrot(mp, f3(jq[JQ_AXIS_P]));
  V3 ref_p = mrot(mp, f3(jq[JQ_AXIS_P + 2]));
  V3 ref_c = mrot(mc, f3(jq[JQ_AXIS_C + 2]));
  axis[0] = ax;
  angle[0] = signed_angle(ax, ref_p, ref_c);
}

// Spherical.axis_angle joints.py:388-415 (spring_joints.py:212-224, 276-303)
BXS void axis_angle_spherical(const float4* jq, const M3& mp, const M3& mc, V3* axis, float* angle) {
  V3 a1p = mrot(mp, f3(jq[JQ_AXIS_P]));
  V3 a2p = mrot(mp, f3(jq[JQ_AXIS_P + 1]));
  V3 a1c = mrot(mc, f3(jq[JQ_AXIS_C]));
  V3 a2c = mrot(mc, f3(jq[JQ_AXIS_C + 1]));
  V3 a3c = mrot(mc, f3(jq[JQ_AXIS_C + 2]));
  V3 lon = cross(a3c, a1p);
  lon = vdiv(lon, 1e-10f + safe_norm(lon));
  float psi = signed_angle(a1p, a2p, lon);
  V3 xz = dot(a1p, a1c) * a1c + dot(a1p, a2c) * a2c;
  xz = vdiv(xz, 1e-10f + safe_norm(xz));
  float theta = bx_acos(clipf(dot(xz, a1p), -1.f, 1.f)) * signf(dot(a1p, a3c));
  float phi = signed_angle(-a3c, a2c, lon);
  axis[0] = a1p; axis[1] = a2c; axis[2] = a3c;
  angle[0] = psi; angle[1] = theta; angle[2] = phi;
}

// The same axes, and per dof only WHETHER the angle lies outside [lo, hi]: all a Torque actuator asks
// of the angle (actuators.py:103-104).  psi / phi are signed angles: decided from their (cos, sin)
// without arctan2 (bx_math.cuh angle_outside); theta is compared as a value.
BXS void axis_outside_revolute(const float4* jq, const M3& mp, const M3& mc, V3* axis, bool* outside) {
  V3 ax = mrot(mp, f3(jq[JQ_AXIS_P]));
  V3 ref_p = mrot(mp, f3(jq[JQ_AXIS_P + 2]));
  V3 ref_c = mrot(mc, f3(jq[JQ_AXIS_C + 2]));
  const float4 l01 = jq[JQ_LIMIT];
  bool below, above;
  angle_outside(dot(ref_p, ref_c), dot(cross(ref_p, ref_c), ax), limcs(jq, 0), l01.x, l01.y, &below, &above);
  axis[0] = ax;
  outside[0] = below || above;
}
BXS void axis_outside_spherical(const float4* jq, const M3& mp, const M3& mc, V3* axis, bool* outside) {
  V3 a1p = mrot(mp, f3(jq[JQ_AXIS_P]));
  V3 a2p = mrot(mp, f3(jq[JQ_AXIS_P + 1]));
  V3 a1c = mrot(mc, f3(jq[JQ_AXIS_C]));
  V3 a2c = mrot(mc, f3(jq[JQ_AXIS_C + 1]));
  V3 a3c = mrot(mc, f3(jq[JQ_AXIS_C + 2]));
  const float4 l01 = jq[JQ_LIMIT], l2 = jq[JQ_LIMIT + 1];
  V3 lon = cross(a3c, a1p);
  lon = vdiv(lon, 1e-10f + safe_norm(lon));
  bool below, above;
  angle_outside(dot(a2p, lon), dot(cross(a2p, lon), a1p), limcs(jq, 0), l01.x, l01.y, &below, &above);
  outside[0] = below || above;
  V3 xz = dot(a1p, a1c) * a1c + dot(a1p, a2c) * a2c;
  xz = vdiv(xz, 1e-10f + safe_norm(xz));
  float theta = bx_acos(clipf(dot(xz, a1p), -1.f, 1.f)) * signf(dot(a1p, a3c));
  outside[1] = theta < l01.z || theta > l01.w;
  angle_outside(dot(a2c, lon), dot(cross(a2c, lon), -a3c), limcs(jq, 2), l2.x, l2.y, &below, &above);
  outside[2] = below || above;
  axis[0] = a1p; axis[1] = a2c; axis[2] = a3c;
}
BXS void joint_axis_outside(int kind, const float4* jq, const M3& mp, const M3& mc, V3* axis, bool* outside) {
  if (BXF_IS(kind == BX_JOINT_REVOLUTE || kind == BX_JOINT_SPRING_REVOLUTE, BXF_REVOLUTE,
             BXF_SPHERICAL | BXF_UNIVERSAL))
    axis_outside_revolute(jq, mp, mc, axis, outside);
  else
    axis_outside_spherical(jq, mp, mc, axis, outside);
}

BXS int joint_dof(int kind) {
  return BXF_IS(kind == BX_JOINT_REVOLUTE || kind == BX_JOINT_SPRING_REVOLUTE, BXF_REVOLUTE,
                BXF_SPHERICAL | BXF_UNIVERSAL)
             ? 1
             : (BXF_IS(kind == BX_JOINT_SPRING_UNIVERSAL, BXF_UNIVERSAL, BXF_SPHERICAL) ? 2 : 3);
}

BXS void joint_axis_angle(int kind, const float4* jq, const M3& mp, const M3& mc, V3* axis, float* angle) {
  if (BXF_IS(kind == BX_JOINT_REVOLUTE || kind == BX_JOINT_SPRING_REVOLUTE, BXF_REVOLUTE,
             BXF_SPHERICAL | BXF_UNIVERSAL))
    axis_angle_revolute(jq, mp, mc, axis, angle);
  else
    axis_angle_spherical(jq, mp, mc, axis, angle);
}
BXS void joint_axis_angle(int kind, const float4* jq, Q4 rp, Q4 rc, V3* axis, float* angle) {
  joint_axis_angle(kind, jq, rotmat(rp), rotmat(rc), axis, angle);
}

// ---- phase T: actuator torques and joint damping torques -------------------
// Writes the torque VECTOR of actuator a to quad JS+a and the damping torque of
// joint j to quad JS+na+j; the body phase turns them into angular
// accelerations (actuators.py:52-112, joints.py:103-128).
template <int MODE>
BXS void phase_torques(const Ctx& c, const BxHeader& h, bool damp) {
  typedef Lay<MODE> L;
  const int ro = MODE == BX_MODE_PBD ? 2 * c.cur + 1 : 1;
  for (int a = c.warp; a < h.na; a += c.nwarp) {
    const int4 ai = ldi4(c.ti + h.io_act + a * AI_STRIDE);      // kind, joint, idx0, idx1
    const int4 ai2 = ldi4(c.ti + h.io_act + a * AI_STRIDE + 4);  // idx2, slot_p, slot_c
    const int idx2 = ai2.x;
    const int j = ai.y;
    const int4 ji = ldi4(c.ti + h.io_joint + j * JI_STRIDE);
    const float4* jq = c.tf + h.fo_joint + j * JQS;
    const int kind = ji.x;
    Q4 rp = ldq(c, L::BQ * ji.y + ro);
    Q4 rc = ldq(c, L::BQ * ji.z + ro);
    V3 axis[3];
    float angle[3] = {0.f, 0.f, 0.f};
    bool outside[3] = {false, false, false};
    const bool torque_act = BXF_IS(ai.x == BX_ACT_TORQUE, BXF_TORQUE, BXF_ANGLE);
    {
      const M3 Rp = rotmat(rp), Rc = rotmat(rc);
      if (BX_TORQUE_ARC && torque_act) joint_axis_outside(kind, jq, Rp, Rc, axis, outside);
      else joint_axis_angle(kind, jq, Rp, Rc, axis, angle);
    }
    const float strength = c.tf[h.fo_act + a].x;
    const int dof = joint_dof(kind);
    const float4 l01 = jq[JQ_LIMIT], l2 = jq[JQ_LIMIT + 1];
    const float lo[3] = {l01.x, l01.z, l2.x}, hi[3] = {l01.y, l01.w, l2.y};
    const int aidx[3] = {ai.z, ai.w, idx2};
    V3 tq = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      if (d < dof) {
        int idx = aidx[d];
        // take(act, idx, mode=clip) * act_mask  (actuators.py:64)
        float m = idx >= 0 ? 1.f : 0.f;
        idx = min(max(idx, 0), h.act_size - 1);
        float av = act_at(c, h, idx) * m;
        float t;
        if (torque_act) {
          t = av * strength * -1.f;
          t = (BX_TORQUE_ARC ? outside[d] : (angle[d] < lo[d] || angle[d] > hi[d])) ? 0.f : t;   // actuators.py:103-104
        } else {
          float tgt = clipf(av * 0.017453292519943295f, lo[d], hi[d]);  // act * pi / 180 (actuators.py:83)
          t = (tgt - angle[d]) * strength;
        }
        tq = tq + axis[d] * t;
      }
    }
    // signed I^-1 tau straight into the two bodies' actuator lists.  Torque:
    // parent +, child -; Angle: parent -, child + (actuators.py:88-89,109-110)
    V3 tp = f3(c.tf[h.fo_body + ji.y * BQ_STRIDE + BQ_INV_INERTIA]) * tq;
    V3 tc = f3(c.tf[h.fo_body + ji.z * BQ_STRIDE + BQ_INV_INERTIA]) * tq;
    const bool angle_act = BXF_IS(ai.x == BX_ACT_ANGLE, BXF_ANGLE, BXF_TORQUE);
    stv(c, h.q_js + ai2.y, angle_act ? -tp : tp);
    stv(c, h.q_js + ai2.z, angle_act ? tc : -tc);
  }
  if (damp) {
    for (int j = c.damp_j0; j < h.nj; j += c.damp_step) {
      const int4 ji = ldi4(c.ti + h.io_joint + j * JI_STRIDE);
      const int4 js = ldi4(c.ti + h.io_joint + j * JI_STRIDE + 4);  // dslot_p dslot_c jslot_p jslot_c
      const float damping = c.tf[h.fo_joint + j * JQS + JQ_OFF_P].w;
      V3 ap = ldv(c, L::BQ * ji.y + L::ANG);
      V3 ac = ldv(c, L::BQ * ji.z + L::ANG);
      V3 tq = (-1.f * damping) * (ap - ac);
      stv(c, h.q_js + js.x, f3(c.tf[h.fo_body + ji.y * BQ_STRIDE + BQ_INV_INERTIA]) * tq);
      stv(c, h.q_js + js.y, -(f3(c.tf[h.fo_body + ji.z * BQ_STRIDE + BQ_INV_INERTIA]) * tq));
    }
  }
}

// Sum of the angular accelerations of body b from actuators (dp_a), joint
// damping (dp_j) and forces (dp_f); system.py:268-271.
BXS void body_accels(const Ctx& c, const BxHeader& h, const int32_t* bi, const float4* bq,
                     bool damp, V3* dpa_ang, V3* dpj_ang, V3* dpf_vel, V3* dpf_ang) {
  V3 dpa = v3(0.f, 0.f, 0.f), dpj = v3(0.f, 0.f, 0.f);
  {
    const int4 b0 = ldi4(bi);  // cslot aslot a_base a_n
    const int r = h.q_js + b0.z;
#pragma unroll 1
    for (int i = 0; i < b0.w; i++) dpa = dpa + ldv(c, r + i);
  }
  if (damp) {
    const int r = h.q_js + bi[BI_D_BASE], n = bi[BI_D_N];
#pragma unroll 1
    for (int i = 0; i < n; i++) dpj = dpj + ldv(c, r + i);
  }
  V3 fv = v3(0.f, 0.f, 0.f), fa = v3(0.f, 0.f, 0.f);
  if (BX_FEAT & BXF_FORCES) {
    const int32_t* adj = c.ti + h.io_adj + bi[BI_F_OFF];
    const int n = bi[BI_F_N];
#pragma unroll 1
    for (int i = 0; i < n; i++) {
      int f = adj[i];
      const int32_t* fi = c.ti + h.io_force + f * FI_STRIDE;
      float strength = c.tf[h.fo_force + f].x;
      V3 a = v3(act_at(c, h, min(max(fi[FI_IDX], 0), h.act_size - 1)),
                act_at(c, h, min(max(fi[FI_IDX + 1], 0), h.act_size - 1)),
                act_at(c, h, min(max(fi[FI_IDX + 2], 0), h.act_size - 1)));
      V3 d = vdiv(a * strength, bq[BQ_MASS].y);  // forces.py:44,87
      if (fi[FI_KIND] == BX_FORCE_THRUSTER) fv = fv + d; else fa = fa + d;
    }
  }
  *dpa_ang = dpa; *dpj_ang = dpj; *dpf_vel = fv; *dpf_ang = fa;
}

// ---- integrator pieces (per body) -------------------------------------------
// Euler.kinetic integrators.py:50-68
BXS void kinetic(const BxHeader& h, V3 pm, V3 rm, V3 vel, V3 ang, V3* pos, Q4* rot) {
  *pos = *pos + (vel * h.dt) * pm;
  V3 w = (ang * rm) * h.half_dt;
  Q4 r = *rot + quat_mul(q4(0.f, w.x, w.y, w.z), *rot);
#if BX_FAST_DIV && BX_NORM_RSQRT
  float rn = bx_rsqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);  // rot / norm(rot)
  *rot = q4(r.w * rn, r.x * rn, r.y * rn, r.z * rn);
#else
  float n = bx_sqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
  float rn = bx_rcp(n);
  *rot = q4(bx_mul_rcp(r.w, n, rn), bx_mul_rcp(r.x, n, rn), bx_mul_rcp(r.y, n, rn), bx_mul_rcp(r.z, n, rn));
#endif
}

// Euler.velocity_projection integrators.py:122-146
BXS void velocity_projection(const BxHeader& h, V3 pm, V3 rm, V3 pos, V3 ppos, Q4 prot, Q4* rot,
                             V3* vel, V3* ang) {
  Q4 r = *rot;
#if BX_FAST_DIV && BX_NORM_RSQRT
  float rn = bx_rsqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
  Q4 nr = q4(r.w * rn, r.x * rn, r.y * rn, r.z * rn);
#else
  float n = bx_sqrt(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
  float rn = bx_rcp(n);
  Q4 nr = q4(bx_mul_rcp(r.w, n, rn), bx_mul_rcp(r.x, n, rn), bx_mul_rcp(r.y, n, rn), bx_mul_rcp(r.z, n, rn));
#endif
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
BXS void phase_body_acc_kinetic(const Ctx& c, const BxHeader& h, bool accumulate_info) {
  typedef Lay<BX_MODE_PBD> L;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    PTM0(c);
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b;
    V3 dpa, dpj, fv, fa;
    body_accels(c, h, bi, bq, true, &dpa, &dpj, &fv, &fa);
    V3 acc_ang = (dpa + fa) + dpj;
    PTM(c, 0);
    V3 vel = ldv(c, r0 + L::VEL), ang = ldv(c, r0 + L::ANG);
    V3 g = v3(h.gravity[0], h.gravity[1], h.gravity[2]);
    V3 pm = f3(bq[BQ_POS_MASK]), rm = f3(bq[BQ_ROT_MASK]);
    // Euler.update acc branch integrators.py:87-95
    vel = (h.vel_damp * vel + (fv + g) * h.dt) * pm;
    ang = (h.ang_damp * ang + acc_ang * h.dt) * rm;
    V3 pos = ldv(c, r0 + 2 * c.cur);
    Q4 rot = ldq(c, r0 + 2 * c.cur + 1);
    PTM(c, 1);
    kinetic(h, pm, rm, vel, ang, &pos, &rot);
    PTM(c, 2);
    stv(c, r0 + L::VEL, vel);
    stv(c, r0 + L::ANG, ang);
    stv(c, r0 + 2 * (c.cur ^ 1), pos);
    stq(c, r0 + 2 * (c.cur ^ 1) + 1, rot);
    PTM(c, 3);
    if (accumulate_info && bi[BI_ASLOT] >= 0) {  // info.actuator += dp_a (system.py:315)
      int r = h.q_ia + bi[BI_ASLOT];
      stv(c, r, ldv(c, r) + dpa);
    }
  }
}

// ---- pbd joints: joints.py ---------------------------------------------------
struct DQ { V3 pos; Q4 rot; };

// Joint.apply_angle_update joints.py:130-152 (rot parts only; pos parts are 0)
BXS void angle_update(V3 dq, V3 ii_p, V3 ii_c, Q4 rp, Q4 rc, float scale_ang,
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
BXS void position_update(V3 wp, V3 wc, V3 pos_p, V3 pos_c, Q4 rp, Q4 rc,
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
  dp->pos = scale_pos * vdivm(p, m_p, im_p);
  dp->rot = scale_pos * (.5f * vec_quat_mul(ii_p * cross(r_p, p), rp));
  dc->pos = scale_pos * vdivm(-p, m_c, im_c);
  dc->rot = scale_pos * (-.5f * vec_quat_mul(ii_c * cross(r_c, p), rc));
}

// limit_angle closure of Spherical.apply_reduced joints.py:343-354
BXS void limit_angle(V3 n, V3 n_1, V3 n_2, float lo, float hi, float4 lcs, V3 ii_p, V3 ii_c,
                     Q4 rp, Q4 rc, float scale_ang, Q4* out_p, Q4* out_c) {
  // ph = signed_angle(n, n_1, n_2); mask = ph outside [lo, hi]; n1 = n_1 rotated about n by clip(ph)
  float cs, sn, mask;
  clip_dir(dot(n_1, n_2), dot(cross(n_1, n_2), n), lcs, lo, hi, &cs, &sn, &mask);
  V3 n1 = rot_axis_cs(n_1, n, cs, sn);
  V3 dq = cross(n1, n_2) * mask;
  angle_update(dq, ii_p, ii_c, rp, rc, scale_ang, out_p, out_c);
}

// ---- phase J (pbd): Joint.apply for every joint -> per-joint delta slots -------
// joints.py:79-100, Revolute :270-309, Spherical :332-386.  Slot of joint j:
// quads JS+4j .. +3 = parent dpos, parent drot, child dpos, child drot.
// BXF_JOINT_OFFSET builds: the joint sockets come from this env's own table (bx_step_io.joint_offset).  A build of
// its own (all features, bx_kernels.cu) because every cheaper arrangement taxed all systems -- measured: a per-item
// test of h.jo_off 1 %, a second inlined instantiation picked per launch 0.9 % (the step loop's code grows by a phase
// that never runs: instruction cache), the same out of line 0.5-1 % (the call constrains register allocation).
BXS void phase_joints_pbd(const Ctx& c, const BxHeader& h) {
  typedef Lay<BX_MODE_PBD> L;
  for (int j = c.warp; j < h.nj; j += c.nwarp) {
    PTM0(c);
    const int4 ji = ldi4(c.ti + h.io_joint + j * JI_STRIDE);
    const float4* jq = c.tf + h.fo_joint + j * JQS;
    const int bp = ji.y, bc = ji.z;
    const float4* bqp = c.tf + h.fo_body + bp * BQ_STRIDE;
    const float4* bqc = c.tf + h.fo_body + bc * BQ_STRIDE;
    V3 pos_p = ldv(c, L::BQ * bp + 2 * c.cur), pos_c = ldv(c, L::BQ * bc + 2 * c.cur);
    Q4 rp = ldq(c, L::BQ * bp + 2 * c.cur + 1), rc = ldq(c, L::BQ * bc + 2 * c.cur + 1);
    V3 ii_p = f3(bqp[BQ_INV_INERTIA]), ii_c = f3(bqc[BQ_INV_INERTIA]);
    const float4 mp = bqp[BQ_MASS], mc = bqc[BQ_MASS];
    float4 j0 = jq[JQ_OFF_P];
    const float4 j1 = jq[JQ_OFF_C];
    if constexpr ((BX_FEAT & BXF_JOINT_OFFSET) != 0) {   // this env's own socket (bx_step_io.joint_offset)
      const float4 o = bx_smem[h.jo_off + j * BX_ENVS_PER_CTA + (threadIdx.x & 31)];
      j0.x = o.x; j0.y = o.y; j0.z = o.z;
    }
    const float4 ap0 = jq[JQ_AXIS_P];
    const M3 Rp = rotmat(rp), Rc = rotmat(rc);
    V3 wp = pos_p + mrot(Rp, f3(j0));
    V3 wc = pos_c + mrot(Rc, f3(j1));
    PTM(c, 4);
    DQ dp, dc;
    position_update(wp, wc, pos_p, pos_c, rp, rc, mp.x, mc.x, mp.y, mc.y, ii_p, ii_c, j1.w, &dp, &dc);
    PTM(c, 5);
    const float sa = ap0.w;
    const float4 l01 = jq[JQ_LIMIT];
    if (BXF_IS(ji.x == BX_JOINT_REVOLUTE, BXF_REVOLUTE, BXF_SPHERICAL)) {
      V3 axis = mrot(Rp, f3(ap0));
      V3 ref_p = mrot(Rp, f3(jq[JQ_AXIS_P + 2]));
      V3 ref_c = mrot(Rc, f3(jq[JQ_AXIS_C + 2]));
      // psi = signed_angle(axis, ref_p, ref_c); n1 = ref_p rotated about axis by clip(psi, lo, hi)
      float cs, sn, outside;
      clip_dir(dot(ref_p, ref_c), dot(cross(ref_p, ref_c), axis), limcs(jq, 0), l01.x, l01.y, &cs, &sn, &outside);
      PTM(c, 6);
      V3 axis_c = mrot(Rc, f3(jq[JQ_AXIS_C]));
      V3 dq_1 = cross(axis, axis_c);
      V3 n1 = rot_axis_cs(ref_p, axis, cs, sn);
      V3 dq_2 = cross(n1, ref_c);
      PTM(c, 7);
      Q4 p1, c1, p2, c2;
      angle_update(dq_1, ii_p, ii_c, rp, rc, sa, &p1, &c1);
      angle_update(dq_2, ii_p, ii_c, rp, rc, sa, &p2, &c2);
      dp.rot = dp.rot + (p1 + p2);
      dc.rot = dc.rot + (c1 + c2);
      PTM(c, 8);
    } else {
      const float4 l2 = jq[JQ_LIMIT + 1];
      V3 a1p = mrot(Rp, f3(ap0));
      V3 a2p = mrot(Rp, f3(jq[JQ_AXIS_P + 1]));
      V3 a1c = mrot(Rc, f3(jq[JQ_AXIS_C]));
      V3 a2c = mrot(Rc, f3(jq[JQ_AXIS_C + 1]));
      V3 a3c = mrot(Rc, f3(jq[JQ_AXIS_C + 2]));
      V3 lon = cross(a3c, a1p);
      lon = vdiv(lon, 1e-6f + safe_norm(lon));
      V3 xz = dot(a1p, a1c) * a1c + dot(a1p, a2c) * a2c;
      xz = vdiv(xz, 1e-6f + safe_norm(xz));
      V3 a2n = cross(xz, a1p);
      a2n = vdiv(a2n, 1e-6f + safe_norm(a2n));
      float sg = signf(dot(a1p, a3c));
      Q4 p1, c1, p2, c2, p3, c3;
      limit_angle(a1p, a2p, lon, l01.x, l01.y, limcs(jq, 0), ii_p, ii_c, rp, rc, sa, &p1, &c1);
      limit_angle((-a2n) * sg, a1p, xz, l01.z, l01.w, limcs(jq, 1), ii_p, ii_c, rp, rc, sa, &p2, &c2);
      limit_angle(a3c, lon, a2c, l2.x, l2.y, limcs(jq, 2), ii_p, ii_c, rp, rc, sa, &p3, &c3);
      dp.rot = dp.rot + ((p1 + p2) + p3);
      dc.rot = dc.rot + ((c1 + c2) + c3);
    }
    const int4 js = ldi4(c.ti + h.io_joint + j * JI_STRIDE + 4);  // dslot_p dslot_c jslot_p jslot_c
    stv(c, h.q_js + js.z, dp.pos); stq(c, h.q_js + js.z + 1, dp.rot);
    stv(c, h.q_js + js.w, dc.pos); stq(c, h.q_js + js.w + 1, dc.rot);
    PTM(c, 9);
  }
}

// ---- phase B2 (pbd): apply joint deltas (+ velocity projection in half A) ---
// integrators.py:103-110 then :122-146; system.py:277-281 / :296-297.
BXS void phase_body_joint_update(const Ctx& c, const BxHeader& h, bool project) {
  typedef Lay<BX_MODE_PBD> L;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b, rc_ = r0 + 2 * c.cur, rp_ = r0 + 2 * (c.cur ^ 1);
    V3 dpos = v3(0.f, 0.f, 0.f);
    Q4 drot = q4(0.f, 0.f, 0.f, 0.f);
    const int jr = h.q_js + bi[BI_J_BASE], jn = bi[BI_J_N];
#pragma unroll 1
    for (int i = 0; i < jn; i++) {
      dpos = dpos + ldv(c, jr + 2 * i);
      drot = drot + ldq(c, jr + 2 * i + 1);
    }
    V3 pm = f3(bq[BQ_POS_MASK]), rm = f3(bq[BQ_ROT_MASK]);
    V3 pos = ldv(c, rc_) + dpos * pm;
    Q4 rot = ldq(c, rc_ + 1);
    rot = q4(rot.w + drot.w, rot.x + drot.x * rm.x, rot.y + drot.y * rm.y, rot.z + drot.z * rm.z);
    if (project) {
      V3 vel, ang;
      velocity_projection(h, pm, rm, pos, ldv(c, rp_), ldq(c, rp_ + 1), &rot, &vel, &ang);
      stv(c, r0 + L::VEL, vel);
      stv(c, r0 + L::ANG, ang);
    }
    stv(c, rc_, pos);
    stq(c, rc_ + 1, rot);
  }
}

// ---- contacts -----------------------------------------------------------------
struct Contact { V3 pos, vel, normal; float pen; };

// One contact ITEM = 1 or 2 consecutive rows of one collider pair (bx_tables.h).  Its
// rows add their per-body deltas here in row order; the item then writes ONE entry per
// side into the body's slot list: {linear xyz, rows with a non-zero linear part},
// {second part}.  (The row count replaces the `any(x != 0)` test of the gather,
// colliders.py:187-196.)
struct CAcc {
  float4 la, sa, lb, sb;
};
BXS CAcc cacc_zero() {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  CAcc a; a.la = z; a.sa = z; a.lb = z; a.sb = z;
  return a;
}
BXS void cacc_add(float4& l, float4& s2, V3 lin, float4 sec) {
  l.x += lin.x; l.y += lin.y; l.z += lin.z;
  l.w += (lin.x != 0.f || lin.y != 0.f || lin.z != 0.f) ? 1.f : 0.f;
  s2.x += sec.x; s2.y += sec.y; s2.z += sec.z; s2.w += sec.w;
}
BXS float4 wxyz(Q4 q) { return make_float4(q.w, q.x, q.y, q.z); }
BXS float4 xyz0(V3 v) { return make_float4(v.x, v.y, v.z, 0.f); }
// scalar k of a packed scalar region (4 per quad row)
BXS float& scalar_at(const Ctx& c, int row0, int k) {
  return reinterpret_cast<float*>(&QROW(c, row0 + (k >> 2)))[k & 3];
}

// geometry.py:360-374
BXS V3 closest_segment_point_and_dist(V3 a, V3 b, V3 pt, float* d2) {
  V3 ab = b - a;
  float t = bx_div(dot(pt - a, ab), dot(ab, ab) + 1e-6f);
  V3 p = a + clipf(t, 0.f, 1.f) * ab;
  V3 v = pt - p;
  *d2 = dot(v, v);
  return p;
}

// geometry.py:394-459 closest_segment_to_segment_points
BXS void closest_seg_seg(V3 a0, V3 a1, V3 b0, V3 b1, V3* best_a, V3* best_b) {
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

// geometry.py:377-391
BXS V3 closest_segment_point_plane(V3 a, V3 b, V3 p0, V3 n) {
  const V3 ab = b - a;
  const float t = bx_div(dot(p0, n) - dot(n, a), dot(n, ab) + 1e-6f);
  return a + clipf(t, 0.f, 1.f) * ab;
}
// geometry.py:462-500
BXS V3 closest_triangle_point(V3 p0, V3 p1, V3 p2, V3 pt) {
  const V3 e0 = p1 - p0, e1 = p2 - p0, d = pt - p0;
  const float a = dot(e0, e0), b = dot(e0, e1), cc = dot(e1, e1);
  const float rdet = bx_rcp(a * cc - b * b);
  const float u = (cc * dot(e0, d) - b * dot(e1, d)) * rdet;
  const float v = (-b * dot(e0, d) + a * dot(e1, d)) * rdet;
  const bool inside = 0.f <= u && u <= 1.f && 0.f <= v && v <= 1.f && u + v <= 1.f;
  V3 cp = p0 + u * e0 + v * e1;
  const V3 df = cp - pt;
  const float d0 = dot(df, df);
  float d1, d2, d3;
  const V3 c1 = closest_segment_point_and_dist(p0, p1, pt, &d1);
  const bool take0 = d0 < d1 && inside;
  cp = take0 ? cp : c1;
  float min_d = take0 ? d0 : d1;
  const V3 c2 = closest_segment_point_and_dist(p1, p2, pt, &d2);
  cp = d2 < min_d ? c2 : cp;
  min_d = fminf(min_d, d2);
  const V3 c3 = closest_segment_point_and_dist(p2, p0, pt, &d3);
  return d3 < min_d ? c3 : cp;
}
// geometry.py:503-540: the closest of the three edge candidates and the in-triangle candidate;
// candidates that tie are averaged (mask = distance == min; sum * mask / sum(mask))
BXS void closest_segment_triangle_points(V3 a, V3 b, V3 p0, V3 p1, V3 p2, V3 tn, V3* seg_pt, V3* tri_pt) {
  V3 sp[4], tp[4];
  float dd[4];
  closest_seg_seg(a, b, p0, p1, &sp[0], &tp[0]);
  closest_seg_seg(a, b, p1, p2, &sp[1], &tp[1]);
  closest_seg_seg(a, b, p0, p2, &sp[2], &tp[2]);
  sp[3] = closest_segment_point_plane(a, b, p0, tn);
  tp[3] = closest_triangle_point(p0, p1, p2, sp[3]);
#pragma unroll
  for (int i = 0; i < 4; i++) { const V3 v = sp[i] - tp[i]; dd[i] = dot(v, v); }
  const float mn = fminf(fminf(dd[0], dd[1]), fminf(dd[2], dd[3]));
  V3 ss = v3(0.f, 0.f, 0.f), ts = ss;
  float cnt = 0.f;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float m = dd[i] == mn ? 1.f : 0.f;
    cnt += m;
    ss = ss + sp[i] * m;
    ts = ts + tp[i] * m;
  }
  *seg_pt = vdiv(ss, cnt);
  *tri_pt = vdiv(ts, cnt);
}

BXS float w_of(float inv_mass, V3 inv_i, V3 r, V3 n) {
  V3 cr = cross(r, n);
  return inv_mass + dot(cr, inv_i * cr);
}

// Friction of a collider pair: the pair table's, or with per-env collider friction (bx_step_io.friction:
// domain randomisation, experimental/tracing/randomizers.py:26-51) this env's friction^2 from shared memory.
BXS float pair_friction(const BxHeader& h, float table) {
  return h.fr_off ? reinterpret_cast<const float*>(bx_smem)[h.fr_off + (threadIdx.x & 31)] : table;
}

// contact function of one contact row (colliders.py:667-680 box_plane,
// :744-759 capsule_plane, :805-819 capsule_capsule).
BXS Contact contact_fn(const BxHeader& h, int kind, const float4* cq, const float4* tq, V3 pa, Q4 ra, V3 va, V3 wa, V3 pb,
                       Q4 rb, V3 vb, V3 wb) {
  Contact ct;
  const float4 mat = cq[CQ_MAT];
  if ((BX_FEAT & BXF_CAPSULE_BOX) && kind == BX_PAIR_BOX_HEIGHTMAP) {
    // box_heightmap (colliders.py:699-742): this row's box corner against the triangle of the height grid
    // under it; tq[0] = {cell_size, grid size n, offset of the grid in h.hm}
    const V3 off = rotate(f3(cq[CQ_PTS]), ra);
    ct.pos = pa + off;                  // QP.to_world base.py:112-125
    ct.vel = va + cross(wa, off);
    const Q4 rbi = q4(rb.w, -rb.x, -rb.y, -rb.z);
    const V3 lp = rotate(ct.pos - pb, rbi);   // math.inv_rotate
    const float cell = tq[0].x;
    const int n = static_cast<int>(tq[0].y);
    const float* grid = h.hm + static_cast<int>(tq[0].z);
    const float u = __fdiv_rn(lp.x, cell), v = __fdiv_rn(lp.y, cell);
    const float fu = floorf(u), fv = floorf(v);
    const int iu = static_cast<int>(fu), iv = static_cast<int>(fv);
    const bool lower = (u - fu) + (v - fv) < 1.f;   // the square's two triangles (:716-717)
    const float mu = lower ? -1.f : 1.f;
    const int tu0 = iu + (lower ? 0 : 1), tu1 = iu + (lower ? 1 : 0), tu2 = tu0;
    const int tv0 = iv + (lower ? 0 : 1), tv1 = tv0, tv2 = iv + (lower ? 1 : 0);
    // height[u, -v] (:726): a negative index wraps once, anything still outside is clamped (jax gather)
    auto at = [&](int a, int b) {
      b = -b;
      a = a < 0 ? a + n : a; b = b < 0 ? b + n : b;
      a = min(max(a, 0), n - 1); b = min(max(b, 0), n - 1);
      return __ldg(grid + a * n + b);
    };
    const float h0 = at(tu0, tv0), h1 = at(tu1, tv1), h2 = at(tu2, tv2);
    const V3 raw = v3(mu * (h1 - h0), mu * (h2 - h0), cell);
    const V3 nl = vdiv(raw, safe_norm(raw));
    const V3 p0 = v3(static_cast<float>(tu0) * cell, static_cast<float>(tv0) * cell, h0);
    ct.pen = dot(p0 - lp, nl);
    ct.normal = rotate(nl, rb);
    return ct;
  }
  if ((BX_FEAT & BXF_CAPSULE_BOX) && kind == BX_PAIR_CAPSULE_BOX) {
    // capsule_mesh (colliders.py:822-849) against one triangle of the box: tq = {p0, p1, p2, normal}
    const M3 ma = rotmat(ra), mb = rotmat(rb);
    const V3 ca = pa + mrot(ma, f3(cq[CQ_PTS + 1])), ea = mrot(ma, f3(cq[CQ_PTS]));
    const V3 p0 = pb + mrot(mb, f3(tq[0])), p1 = pb + mrot(mb, f3(tq[1])), p2 = pb + mrot(mb, f3(tq[2]));
    const V3 tn = mrot(mb, f3(tq[3]));
    V3 seg_p, tri_p;
    closest_segment_triangle_points(ca + ea, ca - ea, p0, p1, p2, tn, &seg_p, &tri_p);
    const V3 pv = seg_p - tri_p;
    const float dist = safe_norm(pv);
    ct.normal = vdiv(pv, 1e-6f + dist);
    ct.pen = mat.z - dist;
    ct.pos = tri_p;
    ct.vel = (va + cross(wa, ct.pos - pa)) - (vb + cross(wb, ct.pos - pb));
    return ct;
  }
  if ((BX_FEAT & BXF_CAPSULE_BOX) && kind == BX_PAIR_CAPSULE_CLIPPED_PLANE) {
    // capsule_clippedplane (colliders.py:762-802): tq = {normal, pos, x, y, {halfsize_x, halfsize_y}} of the
    // plane in the body-b frame; this row's capsule end in cq[CQ_PTS]
    const M3 mb = rotmat(rb);
    const V3 n0 = mrot(mb, f3(tq[0]));
    // the normal is oriented so that it points at the capsule body's position (:771-773: qp_a.pos . normal)
    const float dir = dot(pa, n0) > 0.f ? 1.f : -1.f;
    const V3 n = n0 * dir;
    ct.pos = (pa + rotate(f3(cq[CQ_PTS]), ra)) - n * mat.z;
    ct.vel = va + cross(wa, ct.pos - pa);
    const V3 plane_pt = mrot(mb, f3(tq[1])) + pb;
    ct.pen = dot(plane_pt - ct.pos, n);
    ct.normal = n;
    // clip against the four side planes (:779-798): normal * normal_dir is the flipped normal times dir again
    const V3 nx = mrot(mb, f3(tq[2])), ny = mrot(mb, f3(tq[3]));
    const float4 hs = tq[4];
    const V3 nn = n * dir;
    const V3 yn = cross(nn, nx), xn = -cross(nn, ny);
    const V3 rel = ct.pos - plane_pt;
    // point_in_front_of_plane (geometry.py:563-565): (pt - plane_pt) . plane_normal > 1e-6
    const bool open = dot(rel - nx * hs.x, xn) > 1e-6f || dot(rel + nx * hs.x, -xn) > 1e-6f ||
                      dot(rel - ny * hs.y, yn) > 1e-6f || dot(rel + ny * hs.y, -yn) > 1e-6f;
    if (open) ct.pen = -1.f;
    return ct;
  }
  if (BXF_IS(kind == BX_PAIR_CAPSULE_CAPSULE, BXF_CAPSULE_CAPSULE,
             BXF_CAPSULE_PLANE | BXF_BOX_PLANE)) {
    const M3 ma = rotmat(ra), mb = rotmat(rb);
    V3 ca = pa + mrot(ma, f3(cq[CQ_PTS + 1])), ea = mrot(ma, f3(cq[CQ_PTS]));
    V3 cb = pb + mrot(mb, f3(cq[CQ_PTS + 3])), eb = mrot(mb, f3(cq[CQ_PTS + 2]));
    V3 best_a, best_b;
    closest_seg_seg(ca + ea, ca - ea, cb + eb, cb - eb, &best_a, &best_b);
    V3 pv = best_a - best_b;
    float dist = safe_norm(pv);
    ct.normal = vdiv(pv, 1e-6f + dist);
    ct.pen = mat.z + mat.w - dist;
    ct.pos = vdiv(best_a + best_b, 2.f);
    ct.vel = (va + cross(wa, ct.pos - pa)) - (vb + cross(wb, ct.pos - pb));
  } else {
    V3 n = rotate_z(rb);
    V3 off = rotate(f3(cq[CQ_PTS]), ra);
    if (BXF_IS(kind == BX_PAIR_CAPSULE_PLANE, BXF_CAPSULE_PLANE, BXF_BOX_PLANE)) {
      ct.pos = (pa + off) - n * mat.z;
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
// contact in quads CT+2k = {pos, pen}, {normal, dlambda} and its deltas
// (a: dpos, drot [, b: dpos, drot]) in quads JS+delta_off(k).
BXS void c1_row(const Ctx& c, const BxHeader& h, int k, CAcc& acc) {
  typedef Lay<BX_MODE_PBD> L;
  const int co = 2 * c.cur, po = 2 * (c.cur ^ 1);
    const int4 ci = ldi4(c.ti + h.io_contact + k * CI_STRIDE);  // kind two_way a b
    const int pair_row = c.ti[h.io_contact + k * CI_STRIDE + CI_PAIR];
    const float4* cq = c.tf + h.fo_contact + k * CQ_STRIDE;
    const int a = ci.z, b = ci.w;
    const bool two_way = BXF_IS(ci.y != 0, BXF_TWO_WAY, BXF_ONE_WAY);
    const float4* bqa = c.tf + h.fo_body + a * BQ_STRIDE;
    const int ra0 = L::BQ * a, rb0 = L::BQ * b;
    V3 pa = ldv(c, ra0 + co), pb = ldv(c, rb0 + co);
    Q4 ra = ldq(c, ra0 + co + 1), rb = ldq(c, rb0 + co + 1);
    Contact ct = contact_fn(h, ci.x, cq, c.tf + h.fo_tri + c.ti[h.io_contact + k * CI_STRIDE + CI_TRI], pa, ra,
                            ldv(c, ra0 + L::VEL), ldv(c, ra0 + L::ANG), pb, rb,
                            ldv(c, rb0 + L::VEL), ldv(c, rb0 + L::ANG));
    V3 iia = f3(bqa[BQ_INV_INERTIA]);
    const float4 ma4 = bqa[BQ_MASS];
    const float ima = ma4.x, ma = ma4.y;
    const float friction = pair_friction(h, cq[CQ_MAT].x);
    const float cs = h.collide_scale;
    V3 n = ct.normal;
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
      V3 dpos = cs * vdivm(p, ma, ima);
      Q4 drot = cs * (.5f * vec_quat_mul(iia * cross(rp, p), ra));
      // static friction
      V3 r1 = rotate(rp, qconj(ra));
      V3 p1bar = ldv(c, ra0 + po) + rotate(r1, ldq(c, ra0 + po + 1));
      V3 deltap = ct.pos - p1bar;
      V3 dpt = deltap - dot(deltap, n) * n;
      float c2 = safe_norm(dpt);
      V3 n2 = vdiv(dpt, c2 + 1e-6f);
      float w1t = w_of(ima, iia, rp, n2);
      float dlt = bx_div(-c2, w1t + 0.f);
      float sm = fabsf(dlt) < fabsf(friction * dl) ? 1.f : 0.f;
      V3 pt = ((dlt * n2) * sm) * cm;
      dpos = dpos + cs * vdivm(pt, ma, ima);
      drot = drot + cs * (.5f * vec_quat_mul(iia * cross(rp, pt), ra));
      cacc_add(acc.la, acc.sa, dpos, wxyz(drot));
      dl_out = dl * cm;
    } else {
      const float4* bqb = c.tf + h.fo_body + b * BQ_STRIDE;
      V3 iib = f3(bqb[BQ_INV_INERTIA]);
      const float4 mb4 = bqb[BQ_MASS];
      const float imb = mb4.x, mb = mb4.y;
      V3 half = vdiv(n * ct.pen, 2.f);
      V3 rp = (ct.pos - half) - pa, rc = (ct.pos + half) - pb;
      float cc = -ct.pen;
      float w1 = w_of(ima, iia, rp, n), w2 = w_of(imb, iib, rc, n);
      float dl = bx_div(-cc, w1 + w2 + 1e-6f);
      float cm = cc < 0.f ? 1.f : 0.f;
      V3 p = (dl * n) * cm;
      V3 dpa = cs * vdivm(p, ma, ima), dpb = cs * vdivm(-p, mb, imb);
      Q4 dra = cs * (.5f * vec_quat_mul(iia * cross(rp, p), ra));
      Q4 drb = cs * (-.5f * vec_quat_mul(iib * cross(rc, p), rb));
      // static friction
      V3 xa = ct.pos - pa, xb = ct.pos - pb;
      V3 r1 = rotate(xa, qconj(ra)), r2 = rotate(xb, qconj(rb));
      V3 p1bar = ldv(c, ra0 + po) + rotate(r1, ldq(c, ra0 + po + 1));
      V3 p2bar = ldv(c, rb0 + po) + rotate(r2, ldq(c, rb0 + po + 1));
      V3 deltap = (ct.pos - p1bar) - (ct.pos - p2bar);
      V3 dpt = deltap - dot(deltap, n) * n;
      float c2 = safe_norm(dpt);
      V3 n2 = vdiv(dpt, c2 + 1e-6f);
      float w1t = w_of(ima, iia, xa, n2), w2t = w_of(imb, iib, xb, n2);
      float dlt = bx_div(-c2, w1t + w2t);
      float sm = fabsf(dlt) < fabsf(dl) ? 1.f : 0.f;
      V3 pt = ((dlt * n2) * sm) * cm;
      dpa = dpa + cs * vdivm(pt, ma, ima);
      dpb = dpb + cs * vdivm(-pt, mb, imb);
      dra = dra + cs * (.5f * vec_quat_mul(iia * cross(xa, pt), ra));
      drb = drb + cs * (.5f * vec_quat_mul(iib * cross(xb, -pt), rb));
      cacc_add(acc.la, acc.sa, dpa, wxyz(dra));
      cacc_add(acc.lb, acc.sb, dpb, wxyz(drb));
      dl_out = dl;  // un-masked (colliders.py:580)
    }
    stv4(c, h.q_ct + k, ct.pos, ct.pen);
    stv(c, h.q_cn + pair_row, ct.normal);  // the same for every row of the pair
    scalar_at(c, h.q_cd, k) = dl_out;
}
// writes the entry (entries) of one item into the slot lists of its bodies
BXS void cacc_store(const Ctx& c, const BxHeader& h, int k0, const CAcc& acc) {
  const int4 cs4 = ldi4(c.ti + h.io_contact + k0 * CI_STRIDE + 4);  // slot_a slot_b pair .
  QROW(c, h.q_js + cs4.x) = acc.la;
  QROW(c, h.q_js + cs4.x + 1) = acc.sa;
  if (BXF_IS(c.ti[h.io_contact + k0 * CI_STRIDE + CI_TWO_WAY] != 0, BXF_TWO_WAY, BXF_ONE_WAY)) {
    QROW(c, h.q_js + cs4.y) = acc.lb;
    QROW(c, h.q_js + cs4.y + 1) = acc.sb;
  }
}
BXS void c1_item(const Ctx& c, const BxHeader& h, int item) {
  const int2 it = *reinterpret_cast<const int2*>(c.ti + h.io_citem + 2 * item);  // first row, rows
  CAcc acc = cacc_zero();
#pragma unroll 1
  for (int r = 0; r < it.y; r++) c1_row(c, h, it.x + r, acc);
  cacc_store(c, h, it.x, acc);
}
BXS void phase_contacts_position(const Ctx& c, const BxHeader& h) {
  for (int i = c.warp; i < h.nci; i += c.nwarp) c1_item(c, h, i);
}

// Per-body scatter-and-average of the contact rows' deltas, one collider group
// at a time (colliders.py:187-196, 231-240): sum rows, count rows with any
// non-zero linear part, divide by (eps + count), then sum over groups.
// Each side of a row holds two quads: linear part (xyz) and a second part
// (rot wxyz for Q deltas; ang xyz + 0 for P deltas).
BXS void gather_contact_deltas(const Ctx& c, const BxHeader& h, int cg_off, int cg_n, float eps,
                               float4* lin_out, float4* sec_out) {
  float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), so = lo;
  const int32_t* cg = c.ti + h.io_adj + cg_off;
#pragma unroll 1
  for (int g = 0; g < cg_n; g++) {
    const int base = h.q_js + cg[2 * g], n = cg[2 * g + 1];
    float4 ls = make_float4(0.f, 0.f, 0.f, 0.f), ss = ls;
    float cnt = 0.f;
#pragma unroll 1
    for (int i = 0; i < n; i++) {
      const float4 l = QROW(c, base + 2 * i), s = QROW(c, base + 2 * i + 1);
      cnt += l.w;  // rows of this entry with a non-zero linear part
      ls.x += l.x; ls.y += l.y; ls.z += l.z;
      ss.x += s.x; ss.y += s.y; ss.z += s.z; ss.w += s.w;
    }
    const float d = eps + cnt, rd = bx_rcp(d);
    lo.x += bx_mul_rcp(ls.x, d, rd); lo.y += bx_mul_rcp(ls.y, d, rd); lo.z += bx_mul_rcp(ls.z, d, rd);
    so.x += bx_mul_rcp(ss.x, d, rd); so.y += bx_mul_rcp(ss.y, d, rd);
    so.z += bx_mul_rcp(ss.z, d, rd); so.w += bx_mul_rcp(ss.w, d, rd);
  }
  *lin_out = lo; *sec_out = so;
}

// ---- phase B3 (pbd): apply contact deltas, keep qp_right_before's velocities,
// velocity projection.  system.py:299-307.
BXS void b3_body(const Ctx& c, const BxHeader& h, int b) {
  typedef Lay<BX_MODE_PBD> L;
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b, rc_ = r0 + 2 * c.cur, rp_ = r0 + 2 * (c.cur ^ 1);
    const int cslot = bi[BI_CSLOT], cadj_off = bi[BI_CG_OFF], cadj_n = bi[BI_CG_N];
    V3 pm = f3(bq[BQ_POS_MASK]), rm = f3(bq[BQ_ROT_MASK]);
    V3 pos = ldv(c, rc_);
    Q4 rot = ldq(c, rc_ + 1);
    if (cadj_n > 0) {
      float4 dl, dr;  // dr is stored (w,x,y,z)
      gather_contact_deltas(c, h, cadj_off, cadj_n, 1e-6f, &dl, &dr);
      pos = pos + f3(dl) * pm;
      rot = q4(rot.w + dr.x, rot.x + dr.y * rm.x, rot.y + dr.z * rm.y, rot.z + dr.w * rm.z);
      stv(c, rc_, pos);
    }
    if (cslot >= 0) {  // qp_right_before.vel / .ang (system.py:306)
      const int r = h.q_pre + 2 * cslot;
      QROW(c, r) = QROW(c, r0 + L::VEL);
      QROW(c, r + 1) = QROW(c, r0 + L::ANG);
    }
    V3 vel, ang;
    velocity_projection(h, pm, rm, pos, ldv(c, rp_), ldq(c, rp_ + 1), &rot, &vel, &ang);
    stv(c, r0 + L::VEL, vel);
    stv(c, r0 + L::ANG, ang);
    stq(c, rc_ + 1, rot);
}
BXS void phase_body_contact_update(const Ctx& c, const BxHeader& h) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) b3_body(c, h, b);
}

// ---- phase C2 (pbd): _velocity_contact for every contact row ------------------
// colliders.py:155-196, OneWay :379-442, TwoWay :584-658.  Deltas (a: dvel, dang
// [, b: dvel, dang]) go to quads JS+delta_off(k).
BXS void c2_row(const Ctx& c, const BxHeader& h, int k, CAcc& acc) {
  typedef Lay<BX_MODE_PBD> L;
  const int co = 2 * c.cur;
    const int4 ci = ldi4(c.ti + h.io_contact + k * CI_STRIDE);
    const int pair_row = c.ti[h.io_contact + k * CI_STRIDE + CI_PAIR];
    const float4 mat = c.tf[h.fo_contact + k * CQ_STRIDE + CQ_MAT];
    const int a = ci.z, b = ci.w;
    const bool two_way = BXF_IS(ci.y != 0, BXF_TWO_WAY, BXF_ONE_WAY);
    const float4* bqa = c.tf + h.fo_body + a * BQ_STRIDE;
    const int ra0 = L::BQ * a;
    const float4 c0 = QROW(c, h.q_ct + k);
    V3 cpos = f3(c0), n = ldv(c, h.q_cn + pair_row);
    const float pen = c0.w, dlambda = scalar_at(c, h.q_cd, k);
    V3 pa = ldv(c, ra0 + co);
    V3 ra = cpos - pa;
    V3 iia = f3(bqa[BQ_INV_INERTIA]);
    const float4 ma4 = bqa[BQ_MASS];
    const float ima = ma4.x, ma = ma4.y;
    const int rpre_a = h.q_pre + 2 * c.ti[h.io_body + a * BI_STRIDE + BI_CSLOT];
    const float stat = pen > 0.f ? 1.f : 0.f;
    if (!two_way) {
      V3 rel = ldv(c, ra0 + L::VEL) + cross(ldv(c, ra0 + L::ANG), ra);
      float v_n = dot(rel, n);
      V3 v_t = rel - n * v_n;
      float vtn = safe_norm(v_t);
      V3 vdir = vdiv(v_t, 1e-6f + vtn);
      float lim = (pair_friction(h, mat.x) * fabsf(dlambda)) * h.inv_2h;
      V3 dvel = (-vdir) * fminf(lim, vtn);
      V3 angw = cross(ra, vdir);
      float w = ima + dot(angw, angw);  // no inertia here (colliders.py:405-406)
      V3 p_dyn = vdiv(dvel, w + 1e-6f);
      // restitution; qp_right_before has the same pos as now
      V3 rel_old = ldv(c, rpre_a) + cross(ldv(c, rpre_a + 1), ra);
      float v_n_old = dot(rel_old, n);
      V3 dx = n * (-v_n - fminf(mat.y * v_n_old, 0.f));
      float cc = safe_norm(dx);
      V3 n2 = vdiv(dx, cc + 1e-6f);
      float w1 = w_of(ima, iia, ra, n2);
      float dlr = bx_div(cc, w1 + 1e-6f);
      float sink = v_n_old <= -h.vth ? 1.f : 0.f;
      V3 p = (((dlr * n2) * sink) + p_dyn) * stat;
      cacc_add(acc.la, acc.sa, vdivm(p, ma, ima), xyz0(cross(iia * ra, p)));
    } else {
      const float4* bqb = c.tf + h.fo_body + b * BQ_STRIDE;
      const int rb0 = L::BQ * b;
      V3 iib = f3(bqb[BQ_INV_INERTIA]);
      const float4 mb4 = bqb[BQ_MASS];
      const float imb = mb4.x, mb = mb4.y;
      const int rpre_b = h.q_pre + 2 * c.ti[h.io_body + b * BI_STRIDE + BI_CSLOT];
      V3 pb = ldv(c, rb0 + co);
      V3 rb = cpos - pb;
      V3 rel = (ldv(c, ra0 + L::VEL) + cross(ldv(c, ra0 + L::ANG), ra)) -
               (ldv(c, rb0 + L::VEL) + cross(ldv(c, rb0 + L::ANG), rb));
      float v_n = dot(rel, n);
      V3 v_t = rel - n * v_n;
      float vtn = safe_norm(v_t);
      V3 vdir = vdiv(v_t, 1e-6f + vtn);
      float lim = (pair_friction(h, mat.x) * fabsf(dlambda)) * h.inv_2h;
      V3 dvel = (-vdir) * fminf(lim, vtn);
      float w1 = w_of(ima, iia, ra, vdir), w2 = w_of(imb, iib, rb, vdir);
      V3 p_dyn = vdiv(dvel, w1 + w2 + 1e-6f);
      V3 rel_old = (ldv(c, rpre_a) + cross(ldv(c, rpre_a + 1), ra)) -
                   (ldv(c, rpre_b) + cross(ldv(c, rpre_b + 1), rb));
      float v_n_old = dot(rel_old, n);
      V3 dx = n * (-v_n - fminf(mat.y * v_n_old, 0.f));
      V3 rc2 = (cpos + n * pen) - pb;
      float cc = safe_norm(dx);
      V3 n2 = vdiv(dx, cc + 1e-6f);
      float w1r = w_of(ima, iia, ra, n2), w2r = w_of(imb, iib, rc2, n2);
      float dlr = bx_div(cc, w1r + w2r + 1e-6f);
      float sink = v_n_old <= 0.f ? 1.f : 0.f;
      V3 p = (((dlr * n2) * sink) + p_dyn) * stat;
      cacc_add(acc.la, acc.sa, vdivm(p, ma, ima), xyz0(cross(iia * ra, p)));
      cacc_add(acc.lb, acc.sb, vdivm(-p, mb, imb), xyz0(cross(iib * rb, -p)));
    }
}
BXS void c2_item(const Ctx& c, const BxHeader& h, int item) {
  const int2 it = *reinterpret_cast<const int2*>(c.ti + h.io_citem + 2 * item);
  CAcc acc = cacc_zero();
#pragma unroll 1
  for (int r = 0; r < it.y; r++) c2_row(c, h, it.x + r, acc);
  cacc_store(c, h, it.x, acc);
}
BXS void phase_contacts_velocity(const Ctx& c, const BxHeader& h) {
  for (int i = c.warp; i < h.nci; i += c.nwarp) c2_item(c, h, i);
}

// ---- phase B4: velocity-level contact update + Info.contact -----------------
// integrators.py:97-101; system.py:313,315 (pbd) / :365,367 (legacy_spring).
template <int MODE>
BXS void b4_body(const Ctx& c, const BxHeader& h, float eps, int b) {
  typedef Lay<MODE> L;
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const int cadj_n = bi[BI_CG_N];
    if (cadj_n == 0) return;
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b;
    float4 dl, ds;
    gather_contact_deltas(c, h, bi[BI_CG_OFF], cadj_n, eps, &dl, &ds);
    V3 dv = f3(dl), da = f3(ds);
    stv(c, r0 + L::VEL, (ldv(c, r0 + L::VEL) + dv) * f3(bq[BQ_POS_MASK]));
    stv(c, r0 + L::ANG, (ldv(c, r0 + L::ANG) + da) * f3(bq[BQ_ROT_MASK]));
    const int r = h.q_ic + 2 * bi[BI_CSLOT];
    stv(c, r, ldv(c, r) + dv);
    stv(c, r + 1, ldv(c, r + 1) + da);
}
template <int MODE>
BXS void phase_body_velocity_update(const Ctx& c, const BxHeader& h, float eps) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) b4_body<MODE>(c, h, eps, b);
}

// ---- fused contact section (pbd): when every collider is one-way and a body has
// at most two contact rows, a body's rows only touch that body, so C1 -> B3 ->
// C2 -> B4 of one body is a chain inside ONE warp (each lane re-reads only what
// it wrote itself): three barriers and their idle time disappear.  Bodies with
// contacts come first in `forder` so the long items start first.
BXS void phase_contacts_fused(const Ctx& c, const BxHeader& h) {
  // rounds alternate direction, so a second round lands on the warps that had the
  // short (contact-free) bodies in the first
  bool flip = false;
  for (int base = 0; base < h.nb; base += c.nwarp, flip = !flip) {
    const int i = base + (flip ? c.nwarp - 1 - c.warp : c.warp);
    if (i >= h.nb) continue;
    const int b = c.ti[h.io_forder + i];
    const int4 b3 = ldi4(c.ti + h.io_body + b * BI_STRIDE + 12);  // item_off item_n
    const int32_t* rows = c.ti + h.io_adj + b3.x;
#pragma unroll 1
    for (int r = 0; r < b3.y; r++) c1_item(c, h, rows[r]);
    b3_body(c, h, b);
#pragma unroll 1
    for (int r = 0; r < b3.y; r++) c2_item(c, h, rows[r]);
    if (b3.y > 0) b4_body<BX_MODE_PBD>(c, h, 1e-6f, b);
  }
}

// ======================= legacy_spring (system.py:342-377) =====================
// Body.impulse bodies.py:46-59
BXS void body_impulse(float4 bm /* {1/mass, mass} */, V3 inv_i, V3 bpos, V3 imp, V3 at, V3* dvel, V3* dang) {
  *dvel = vdivm(imp, bm.y, bm.x);
  *dang = inv_i * cross(at - bpos, imp);
}

// phase S1: kinetic step for every body, in place (system.py:352).
BXS void phase_body_kinetic(const Ctx& c, const BxHeader& h) {
  typedef Lay<BX_MODE_LEGACY_SPRING> L;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b;
    V3 pos = ldv(c, r0);
    Q4 rot = ldq(c, r0 + 1);
    kinetic(h, f3(bq[BQ_POS_MASK]), f3(bq[BQ_ROT_MASK]), ldv(c, r0 + L::VEL), ldv(c, r0 + L::ANG),
            &pos, &rot);
    stv(c, r0, pos);
    stq(c, r0 + 1, rot);
  }
}

// phase S2: spring joints (spring_joints.py:89-303) -> slot quads
// JS+na+4j: parent dvel, dang, child dvel, dang.
BXS void phase_joints_spring(const Ctx& c, const BxHeader& h) {
  typedef Lay<BX_MODE_LEGACY_SPRING> L;
  for (int j = c.warp; j < h.nj; j += c.nwarp) {
    const int4 ji = ldi4(c.ti + h.io_joint + j * JI_STRIDE);
    const float4* jq = c.tf + h.fo_joint + j * JQS;
    const int kind = ji.x, bp = ji.y, bc = ji.z;
    const float4* bqp = c.tf + h.fo_body + bp * BQ_STRIDE;
    const float4* bqc = c.tf + h.fo_body + bc * BQ_STRIDE;
    const int rp0 = L::BQ * bp, rc0 = L::BQ * bc;
    V3 pos_p = ldv(c, rp0), pos_c = ldv(c, rc0);
    Q4 rp = ldq(c, rp0 + 1), rc = ldq(c, rc0 + 1);
    V3 ang_p = ldv(c, rp0 + L::ANG), ang_c = ldv(c, rc0 + L::ANG);
    V3 ii_p = f3(bqp[BQ_INV_INERTIA]), ii_c = f3(bqc[BQ_INV_INERTIA]);
    float4 j0 = jq[JQ_OFF_P];
    const float4 j1 = jq[JQ_OFF_C];
    if constexpr ((BX_FEAT & BXF_JOINT_OFFSET) != 0) {   // this env's own socket (bx_step_io.joint_offset)
      const float4 o = bx_smem[h.jo_off + j * BX_ENVS_PER_CTA + (threadIdx.x & 31)];
      j0.x = o.x; j0.y = o.y; j0.z = o.z;
    }
    const float stiffness = jq[JQ_AXIS_P + 1].w, spring_damp = jq[JQ_AXIS_P + 2].w;
    const float ang_damp = j0.w, ls = jq[JQ_AXIS_C].w;
    // QP.to_world base.py:112-125
    const M3 mp = rotmat(rp), mc = rotmat(rc);
    V3 op = mrot(mp, f3(j0)), oc = mrot(mc, f3(j1));
    V3 wp = pos_p + op, wc = pos_c + oc;
    V3 vp = ldv(c, rp0 + L::VEL) + cross(ang_p, op);
    V3 vc = ldv(c, rc0 + L::VEL) + cross(ang_c, oc);
    V3 imp = (wp - wc) * stiffness + spring_damp * (vp - vc);
    V3 dvp, dap, dvc, dac;
    body_impulse(bqp[BQ_MASS], ii_p, pos_p, -imp, wp, &dvp, &dap);
    body_impulse(bqc[BQ_MASS], ii_c, pos_c, imp, wc, &dvc, &dac);
    V3 axis[3];
    float angle[3], dang[3];
    joint_axis_angle(kind, jq, mp, mc, axis, angle);
    const int dof = joint_dof(kind);
    const float4 l01 = jq[JQ_LIMIT], l2 = jq[JQ_LIMIT + 1];
    const float lo[3] = {l01.x, l01.z, l2.x}, hi[3] = {l01.y, l01.w, l2.y};
#pragma unroll
    for (int d = 0; d < 3; d++) {
      float v = 0.f;
      if (d < dof) {
        v = angle[d] < lo[d] ? lo[d] - angle[d] : 0.f;
        v = angle[d] > hi[d] ? hi[d] - angle[d] : v;
      }
      dang[d] = v;
    }
    V3 tq;
    if (BXF_IS(kind == BX_JOINT_SPRING_REVOLUTE, BXF_REVOLUTE, BXF_SPHERICAL | BXF_UNIVERSAL)) {  // spring_joints.py:127-148
      V3 axis_c = mrot(mc, f3(jq[JQ_AXIS_C]));
      tq = stiffness * cross(axis[0], axis_c);
      tq = tq - (ls * axis[0]) * dang[0];
    } else if (BXF_IS(kind == BX_JOINT_SPRING_UNIVERSAL, BXF_UNIVERSAL, BXF_SPHERICAL)) {  // :178-210
      V3 proj = axis[1] - dot(axis[1], axis[0]) * axis[0];
      proj = vdiv(proj, safe_norm(proj));
      tq = (ls / 5.f) * cross(proj, axis[1]);
      tq = tq - ls * (axis[0] * dang[0] + axis[1] * dang[1]);
    } else {  // spring Spherical :249-274
      tq = (-ls) * ((axis[0] * dang[0] + axis[1] * dang[1]) + axis[2] * dang[2]);
    }
    tq = tq - ang_damp * (ang_p - ang_c);
    dap = dap + ii_p * tq;
    dac = dac + (-ii_c) * tq;
    const int4 js = ldi4(c.ti + h.io_joint + j * JI_STRIDE + 4);  // - - jslot_p jslot_c
    stv(c, h.q_js + js.z, dvp); stv(c, h.q_js + js.z + 1, dap);
    stv(c, h.q_js + js.w, dvc); stv(c, h.q_js + js.w + 1, dac);
  }
}

// phase S3: acceleration update from joints + actuators + forces; Info.joint and
// Info.actuator accumulation (system.py:356-359,367).
BXS void phase_body_acc_spring(const Ctx& c, const BxHeader& h) {
  typedef Lay<BX_MODE_LEGACY_SPRING> L;
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const float4* bq = c.tf + h.fo_body + b * BQ_STRIDE;
    const int r0 = L::BQ * b;
    V3 dpa, unused, fv, fa;
    body_accels(c, h, bi, bq, false, &dpa, &unused, &fv, &fa);
    V3 jv = v3(0.f, 0.f, 0.f), ja = v3(0.f, 0.f, 0.f);
    const int jr = h.q_js + bi[BI_J_BASE], jn = bi[BI_J_N];
#pragma unroll 1
    for (int i = 0; i < jn; i++) {
      jv = jv + ldv(c, jr + 2 * i);
      ja = ja + ldv(c, jr + 2 * i + 1);
    }
    V3 g = v3(h.gravity[0], h.gravity[1], h.gravity[2]);
    V3 acc_v = jv + fv, acc_a = (ja + dpa) + fa;
    V3 vel = (h.vel_damp * ldv(c, r0 + L::VEL) + (acc_v + g) * h.dt) * f3(bq[BQ_POS_MASK]);
    V3 ang = (h.ang_damp * ldv(c, r0 + L::ANG) + acc_a * h.dt) * f3(bq[BQ_ROT_MASK]);
    stv(c, r0 + L::VEL, vel);
    stv(c, r0 + L::ANG, ang);
    const int rj = h.q_ij + 2 * b;
    stv(c, rj, ldv(c, rj) + jv);
    stv(c, rj + 1, ldv(c, rj + 1) + ja);
    if (bi[BI_ASLOT] >= 0) {
      int r = h.q_ia + bi[BI_ASLOT];
      stv(c, r, ldv(c, r) + dpa);
    }
  }
}

// phase S4: contact_fn + velocity-level _contact (colliders.py:116-153,
// OneWay :267-304, TwoWay :449-493).
// MODE: where the pose rows live (pbd: the current ping-pong buffer) -- the spring-style impulse is
// also what System.info(qp) evaluates in pbd systems (system.py:327-340), see phase_info_contacts.
template <int MODE>
BXS void spring_row(const Ctx& c, const BxHeader& h, int k, CAcc& acc) {
  typedef Lay<MODE> L;
  const int po = MODE == BX_MODE_PBD ? 2 * c.cur : 0;
  {
    const int4 ci = ldi4(c.ti + h.io_contact + k * CI_STRIDE);
    const int pair_row = c.ti[h.io_contact + k * CI_STRIDE + CI_PAIR];
    const float4* cq = c.tf + h.fo_contact + k * CQ_STRIDE;
    const int a = ci.z, b = ci.w;
    const bool two_way = BXF_IS(ci.y != 0, BXF_TWO_WAY, BXF_ONE_WAY);
    const float4* bqa = c.tf + h.fo_body + a * BQ_STRIDE;
    const float4* bqb = c.tf + h.fo_body + b * BQ_STRIDE;
    const int ra0 = L::BQ * a, rb0 = L::BQ * b;
    V3 pa = ldv(c, ra0 + po), pb = ldv(c, rb0 + po);
    Contact ct = contact_fn(h, ci.x, cq, c.tf + h.fo_tri + c.ti[h.io_contact + k * CI_STRIDE + CI_TRI], pa,
                            ldq(c, ra0 + po + 1), ldv(c, ra0 + L::VEL), ldv(c, ra0 + L::ANG),
                            pb, ldq(c, rb0 + po + 1), ldv(c, rb0 + L::VEL), ldv(c, rb0 + L::ANG));
    const float4 mat = cq[CQ_MAT];
    V3 iia = f3(bqa[BQ_INV_INERTIA]), iib = f3(bqb[BQ_INV_INERTIA]);
    const float4 bma = bqa[BQ_MASS], bmb = bqb[BQ_MASS];  // {1/mass (float32 quotient), mass}
    V3 n = ct.normal;
    V3 ra = ct.pos - pa, rb = ct.pos - pb;
    float baum = h.erp * ct.pen;
    float nv = dot(n, ct.vel);
    V3 c1 = cross(iia * cross(ra, n), ra);
    float ang, denom;
    if (two_way) {
      V3 c2 = cross(iib * cross(rb, n), rb);
      ang = dot(n, c1 + c2);
      denom = bma.x + bmb.x + ang;
    } else {
      ang = dot(n, c1);
      denom = bma.x + ang;
    }
    float imp = bx_div(-1.f * (1.f + mat.y) * nv + baum, denom);
    V3 vd = ct.vel - nv * n;
    float nvd = safe_norm(vd);
    float impd = fminf(bx_div(nvd, denom), pair_friction(h, mat.x) * imp);
    V3 dir = vdiv(vd, 1e-6f + nvd);
    float apn = (ct.pen > 0.f && nv < 0.f && imp > 0.f) ? 1.f : 0.f;
    float apd = apn * (nvd > 0.01f ? 1.f : 0.f);
    V3 dvn, dan, dvd, dad;
    body_impulse(bma, iia, pa, imp * n, ct.pos, &dvn, &dan);
    body_impulse(bma, iia, pa, (-impd) * dir, ct.pos, &dvd, &dad);
    cacc_add(acc.la, acc.sa, dvn * apn + dvd * apd, xyz0(dan * apn + dad * apd));
    if (two_way) {
      body_impulse(bmb, iib, pb, (-imp) * n, ct.pos, &dvn, &dan);
      body_impulse(bmb, iib, pb, impd * dir, ct.pos, &dvd, &dad);
      cacc_add(acc.lb, acc.sb, dvn * apn + dvd * apd, xyz0(dan * apn + dad * apd));
    }
    stv4(c, h.q_ct + k, ct.pos, ct.pen);
    stv(c, h.q_cn + pair_row, ct.normal);
  }
}
template <int MODE>
BXS void phase_contacts_spring(const Ctx& c, const BxHeader& h) {
  for (int i = c.warp; i < h.nci; i += c.nwarp) {
    const int2 it = *reinterpret_cast<const int2*>(c.ti + h.io_citem + 2 * i);
    CAcc acc = cacc_zero();
#pragma unroll 1
    for (int r = 0; r < it.y; r++) spring_row<MODE>(c, h, it.x + r, acc);
    cacc_store(c, h, it.x, acc);
  }
}
// System.info(qp).contact (system.py:327-340 `_pbd_info`, :378-390 `_spring_info`): the summed
// Collider.apply(qp) of all colliders -- per body and collider group the spring-style impulses of its
// rows, averaged over the rows with a non-zero impulse (1e-8 + count, colliders.py:149-152) -- into the
// info.contact accumulators only (the state is left alone).  Follows phase_contacts_spring + a barrier.
template <int MODE>
BXS void phase_info_bodies(const Ctx& c, const BxHeader& h) {
  for (int b = c.warp; b < h.nb; b += c.nwarp) {
    const int32_t* bi = c.ti + h.io_body + b * BI_STRIDE;
    const int cadj_n = bi[BI_CG_N];
    if (cadj_n == 0) continue;
    float4 dl, ds;
    gather_contact_deltas(c, h, bi[BI_CG_OFF], cadj_n, 1e-8f, &dl, &ds);
    const int r = h.q_ic + 2 * bi[BI_CSLOT];
    stv(c, r, f3(dl));
    stv(c, r + 1, f3(ds));
  }
}

// ---- optional epilogue: Joint.angle_vel of the new state ----------------------
// joints.py:197-226 before the free_dofs selection (the host mirror applies it):
// angles of the joint's dofs and vel_d = dot(ang_p - ang_c, axis_d), written to
// quads JS+2j = {angle0..2}, JS+2j+1 = {vel0..2} (JS is dead after the last substep).
template <int MODE>
BXS void phase_angle_vel(const Ctx& c, const BxHeader& h) {
  typedef Lay<MODE> L;
  const int ro = MODE == BX_MODE_PBD ? 2 * c.cur + 1 : 1;
  for (int j = c.warp; j < h.nj; j += c.nwarp) {
    const int4 ji = ldi4(c.ti + h.io_joint + j * JI_STRIDE);
    const float4* jq = c.tf + h.fo_joint + j * JQS;
    Q4 rp = ldq(c, L::BQ * ji.y + ro), rc = ldq(c, L::BQ * ji.z + ro);
    V3 axis[3];
    float angle[3];
    joint_axis_angle(ji.x, jq, rp, rc, axis, angle);
    V3 dw = ldv(c, L::BQ * ji.y + L::ANG) - ldv(c, L::BQ * ji.z + L::ANG);
    const int dof = joint_dof(ji.x);
    float a[3], v[3];
#pragma unroll
    for (int d = 0; d < 3; d++) {
      a[d] = d < dof ? angle[d] : 0.f;
      v[d] = d < dof ? dot(dw, axis[d]) : 0.f;
    }
    QROW(c, h.q_js + 2 * j) = make_float4(a[0], a[1], a[2], 0.f);
    QROW(c, h.q_js + 2 * j + 1) = make_float4(v[0], v[1], v[2], 0.f);
  }
}

// ---- optional epilogue: the reference's Env.step after sys.step -----------------
// envs/ant.py:222-281 (Ant.step, _get_obs), envs/humanoid.py:246-340 (Humanoid.step,
// _get_obs, _center_of_mass).  Rows q_env + EQ_* (bx_tables.h) receive everything
// the obs / reward / done / metrics segments point at; the joint angles come from
// phase_angle_vel's rows.  lane = env as everywhere.
template <int MODE>
BXS V3 env_com(const Ctx& c, const BxHeader& h, int po) {  // humanoid.py:332-334
  typedef Lay<MODE> L;
  V3 s = v3(0.f, 0.f, 0.f);
  for (int b = 0; b < h.nb - 1; b++)
    s = s + c.tf[h.fo_body + b * BQ_STRIDE + BQ_MASS].y * ldv(c, L::BQ * b + po);
  return vdiv(s, h.env_msum);
}
// reference point before the step: torso position (ant.py:226) / centre of mass
// (humanoid.py:250); one warp, at the start of every step
BXS float* act_ptr(const Ctx& c, const BxHeader& h, int i) {
  return reinterpret_cast<float*>(&QROW(c, h.q_act + (i >> 2))) + (i & 3);
}
// grasp.py:75: action = min_act + range_act * ((action + 1) / 2), in place
BXS float env_map_act(const Ctx& c, const BxHeader& h, int i) {
  const float* tab = reinterpret_cast<const float*>(c.tf + h.fo_envact);
  float* a = act_ptr(c, h, i);
  const float v = tab[i] + tab[4 * ((h.act_size + 3) >> 2) + i] * ((*a + 1.f) * 0.5f);
  *a = v;
  return v;
}
// Start of every step, ALL warps (ends before the barrier that precedes the first phase).
// ant / humanoid / fetch: the last warp stores the reference point before the step.  grasp
// (grasp.py:75-85): every warp maps its share of the action vector; the last warp maps the
// three palm-target slots itself and moves the (frozen) palm towards the target.
template <int MODE>
static __device__ BX_ENV_INLINE void phase_env_before(const Ctx c, const BxHeader& h) {
  typedef Lay<MODE> L;
  const int po = MODE == BX_MODE_PBD ? 2 * c.cur : 0;
  const bool last = c.warp == c.nwarp - 1;
  if (h.env_kind == BX_ENV_GRASP) {
    const int na = h.act_size - 3;
    for (int i = c.warp; i < na; i += c.nwarp) env_map_act(c, h, i);
    if (last) {
      const V3 target = v3(env_map_act(c, h, na), env_map_act(c, h, na + 1), env_map_act(c, h, na + 2));
      const int row = L::BQ * h.env_ref + po;
      const V3 palm = ldv(c, row);
      const V3 d = target - palm;
      const float nrm = bx_sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
      const float scale = nrm > 2.f ? bx_div(2.f, nrm) : 1.f;   // the hand does not move too fast
      stv(c, row, palm + (scale * d) * 0.15f);
    }
    return;
  }
  if (!last) return;
  V3 ref;
  if (h.env_kind == BX_ENV_HUMANOID) ref = env_com<MODE>(c, h, po);
  else ref = ldv(c, L::BQ * (h.env_kind == BX_ENV_FETCH ? h.env_ref : 0) + po);
  stv(c, h.q_ep, ref);
}
// math.inv_rotate (math.py:43-53): rotate by quat_inv(q) = (w, -x, -y, -z) (math.py:190-199)
BXS V3 inv_rotate(V3 v, Q4 q) { return rotate(v, q4(q.w, -q.x, -q.y, -q.z)); }
BXS float norm3(V3 a) { return bx_sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
// contacts = where(sum(square(info.contact.vel)) > 1e-5, 1, 0) (fetch.py:116-117, grasp.py:176-178)
BXS float env_contact_flag(const Ctx& c, const BxHeader& h, int b) {
  const int cs = c.ti[h.io_body + b * BI_STRIDE + BI_CSLOT];
  if (cs < 0) return 0.f;
  const V3 v = ldv(c, h.q_ic + 2 * cs);
  return (v.x * v.x + v.y * v.y + v.z * v.z) > 0.00001f ? 1.f : 0.f;
}
// Fetch.step / Grasp.step after sys.step (fetch.py:55-122, grasp.py:88-186)
// rand: this tile's [32,3] uniforms of the target teleport for this step, or nullptr
template <int MODE>
static __device__ BX_ENV_INLINE void phase_env_ego(const Ctx c, const BxHeader& h, const float* rand, int nenv) {
  typedef Lay<MODE> L;
  const int po = MODE == BX_MODE_PBD ? 2 * c.cur : 0;
  const V3 pref = ldv(c, L::BQ * h.env_ref + po);
  const Q4 qref = ldq(c, L::BQ * h.env_ref + po + 1);
  // The target body's rows belong to the LAST warp: it is the one that teleports a hit target
  // below (writes that body's pose row), and the obs must see the pose before the teleport
  // (fetch.py:56 vs :88-93, grasp.py:88 vs :131-136) -- no other warp may read that row here.
  const bool last = c.warp == c.nwarp - 1;
  auto local_rows = [&](int b) {
    // pbd: the body's previous-pose rows are dead after the last substep
    const int r0 = MODE == BX_MODE_PBD ? L::BQ * b + 2 * (c.cur ^ 1) : h.q_eloc + 2 * b;
    const V3 pl = inv_rotate(ldv(c, L::BQ * b + po) - pref, qref);
    const V3 vl = inv_rotate(ldv(c, L::BQ * b + L::VEL), qref);
    stv(c, r0, pl);
    stv(c, r0 + 1, vl);
    reinterpret_cast<float*>(&QROW(c, h.q_ecf + (b >> 2)))[b & 3] = env_contact_flag(c, h, b);
  };
  for (int b = c.warp; b < h.nb; b += c.nwarp)
    if (b != h.env_target) local_rows(b);
  if (!last) return;
  local_rows(h.env_target);
  const int E = h.q_env, V = h.q_ev;
  const V3 ptgt = ldv(c, L::BQ * h.env_target + po);
  float hit_any = 0.f;
  if (h.env_kind == BX_ENV_FETCH) {
    const V3 ex = v3(1.f, 0.f, 0.f), ez = v3(0.f, 0.f, 1.f);
    const V3 fwd = rotate(ex, qref), up = rotate(ez, qref);
    const V3 tl = inv_rotate(ptgt - pref, qref);
    const float tlm = norm3(tl);
    stv(c, V + EV_0, fwd);
    stv(c, V + EV_1, up);
    stv4(c, V + EV_2, vdiv(tl, 1e-6f + tlm), tlm);
    const V3 delta = pref - ldv(c, h.q_ep);                 // fetch.py:59
    const V3 rel = ptgt - pref;
    const float dist = norm3(rel);
    const V3 dir = vdiv(rel, 1e-6f + dist);
    const float moving = 0.1f * dot(delta, dir);
    const float is_up = (0.1f * h.env_dt) * dot(up, ez);
    const float height = (0.1f * h.env_dt) * ldv(c, L::BQ * 0 + po).z;   // qp.pos[0, 2] (fetch.py:70)
    const float facing = dot(dir, fwd);
    const float hit = dist < h.env_radius ? 1.f : 0.f;
    const float weighted = hit * facing;
    const float reward = height + moving + is_up + weighted;  // fetch.py:81
    QROW(c, E + EQ_S0) = make_float4(reward, 0.f, hit, weighted);
    QROW(c, E + EQ_S1) = make_float4(moving, is_up, height, 0.f);
    hit_any = hit;
  } else {
    const V3 pobj = ldv(c, L::BQ * h.env_object + po);
    const V3 vobj = ldv(c, L::BQ * h.env_object + L::VEL);
    const V3 hand_vel = ldv(c, L::BQ * h.env_hand + L::VEL);
    const V3 ol = inv_rotate(pobj - pref, qref);
    const float olm = norm3(ol);
    const V3 tl = inv_rotate(ptgt - pref, qref);
    const float tlm = norm3(tl);
    const V3 h2o = pobj - pref;                               // hand_to_obj = object_rel
    const float h2om = norm3(h2o);
    const V3 h2od = vdiv(h2o, 1e-6f + h2om);
    const V3 o2t = ptgt - pobj;                               // obj_to_target = target_rel
    const float o2tm = norm3(o2t);
    const V3 o2td = vdiv(o2t, 1e-6f + o2tm);
    const float heading = dot(h2od, hand_vel);
    const float obj_heading = dot(o2td, vobj);
    stv4(c, V + EV_0, vdiv(ol, 1e-6f + olm), olm);
    stv4(c, V + EV_1, vdiv(tl, 1e-6f + tlm), tlm);
    stv4(c, V + EV_2, h2o, heading);
    stv4(c, V + EV_3, o2td, o2tm);
    QROW(c, V + EV_4) = make_float4(obj_heading, 0.f, 0.f, 0.f);
    const float planar = bx_sqrt(h2o.x * h2o.x + h2o.y * h2o.y + 0.f);   // norm(object_rel * [1,1,0])
    const float k = 0.1f * h.env_dt;
    const float moving_to_object = k * heading;               // dot(hand_vel, object_dir), grasp.py:99
    const float close = bx_div(k * 1.f, 1.f + planar);
    const float moving_to_target = (1.5f * h.env_dt) * obj_heading;
    float touch = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) touch += env_contact_flag(c, h, h.env_touch[i]);
    const float touching = (0.2f * h.env_dt) * touch;
    const float hit = o2tm < h.env_radius ? 1.f : 0.f;
    const float reward = moving_to_object + close + touching + 5.f * hit + moving_to_target;  // grasp.py:121
    QROW(c, E + EQ_S0) = make_float4(reward, 0.f, hit, touching);
    QROW(c, E + EQ_S1) = make_float4(moving_to_object, moving_to_target, close, 0.f);
    hit_any = hit;
  }
  // teleport hit targets (fetch.py:88-93 + :124-134, grasp.py:131-136 + :188-198) with the
  // caller's uniforms: dist = radius + distance * u0, ang = 2 pi u1, z = 1 | height * u2.
  // The obs above is that of the state before the teleport, as in the reference.
  const int lane = threadIdx.x & 31;
  if (rand != nullptr && lane < nenv && hit_any > 0.f) {
    const float u0 = rand[3 * lane], u1 = rand[3 * lane + 1], u2 = rand[3 * lane + 2];
    const float dist = h.env_radius + h.env_tdist * u0;
    float sn, cs;
    sincosf(6.283185307179586f * u1, &sn, &cs);
    stv(c, L::BQ * h.env_target + po, v3(dist * cs, dist * sn, h.env_theight > 0.f ? h.env_theight * u2 : 1.f));
  }
}
template <int MODE>
static __device__ BX_ENV_INLINE void phase_env(const Ctx c, const BxHeader& h) {
  typedef Lay<MODE> L;
  const int po = MODE == BX_MODE_PBD ? 2 * c.cur : 0;
  const int E = h.q_env;
  const bool scalar_warp = c.warp == c.nwarp - 1;
  float ctrl = 0.f, contact = 0.f;
  V3 ref = v3(0.f, 0.f, 0.f);  // torso position (ant) / centre of mass (humanoid) after the step
  if (h.env_kind == BX_ENV_ANT) {
    if (h.env_flags & 1)  // cfrc = clip(info.contact, -1, 1)  (ant.py:271-279)
      for (int b = c.warp; b < h.nb; b += c.nwarp) {
        const int cs = c.ti[h.io_body + b * BI_STRIDE + BI_CSLOT];
        V3 v = v3(0.f, 0.f, 0.f), w = v;
        if (cs >= 0) { v = ldv(c, h.q_ic + 2 * cs); w = ldv(c, h.q_ic + 2 * cs + 1); }
        stv(c, E + EQ_X + 2 * b, v3(clipf(v.x, -1.f, 1.f), clipf(v.y, -1.f, 1.f), clipf(v.z, -1.f, 1.f)));
        stv(c, E + EQ_X + 2 * b + 1, v3(clipf(w.x, -1.f, 1.f), clipf(w.y, -1.f, 1.f), clipf(w.z, -1.f, 1.f)));
      }
    if (scalar_warp) {
      ref = ldv(c, L::BQ * 0 + po);
      for (int b = 0; b < h.nb; b++) {  // contact_cost (ant.py:237-238)
        const int cs = c.ti[h.io_body + b * BI_STRIDE + BI_CSLOT];
        if (cs < 0) continue;
        const V3 v = ldv(c, h.q_ic + 2 * cs);
        const float x = clipf(v.x, -1.f, 1.f), y = clipf(v.y, -1.f, 1.f), z = clipf(v.z, -1.f, 1.f);
        contact += x * x;
        contact += y * y;
        contact += z * z;
      }
      contact *= h.env_wcc;
    }
  } else {
    ref = env_com<MODE>(c, h, po);
    // com_vals (humanoid.py:301-308); Body.inertia is the INVERSE inertia (bodies.py:40)
    for (int b = c.warp; b < h.nb - 1; b += c.nwarp) {
      const float4 bm = c.tf[h.fo_body + b * BQ_STRIDE + BQ_MASS];
      const float4 ii = c.tf[h.fo_body + b * BQ_STRIDE + BQ_INV_INERTIA];
      const V3 d = ldv(c, L::BQ * b + po) - ref;
      const V3 vel = ldv(c, L::BQ * b + L::VEL);
      const float nrm = bx_sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
      const float n2 = nrm * nrm;  // jp.norm(d) ** 2
      const float mn = bm.y * n2;
      const int r0 = E + EQ_X + 5 * b;
      stv(c, r0 + 0, v3(mn + (ii.x - d.x * d.x), -(d.x * d.y), -(d.x * d.z)));
      stv(c, r0 + 1, v3(-(d.y * d.x), mn + (ii.y - d.y * d.y), -(d.y * d.z)));
      stv(c, r0 + 2, v3(-(d.z * d.x), -(d.z * d.y), mn + (ii.z - d.z * d.z)));
      stv(c, r0 + 3, vdiv(bm.y * vel, h.env_msum));
      stv(c, r0 + 4, vdiv(cross(d, vel), 1e-7f + n2));
    }
    // qfrc_actuator = take(action, act_index, clip) * strength (humanoid.py:314-319):
    // masked slots (-1) clamp to action[0], as jp.take does
    for (int i = c.warp; i < h.na; i += c.nwarp) {
      const int32_t* ai = c.ti + h.io_act + i * AI_STRIDE;
      const float st = c.tf[h.fo_act + i].x;
      float q[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int idx = ai[AI_IDX + k];
        q[k] = act_at(c, h, idx < 0 ? 0 : idx) * st;
      }
      stv(c, E + EQ_X + 5 * (h.nb - 1) + i, v3(q[0], q[1], q[2]));
    }
  }
  if (!scalar_warp) return;
  const V3 p0 = ldv(c, L::BQ * 0 + po);
  QROW(c, E + EQ_TORSO_POS) = QROW(c, L::BQ * 0 + po);
  QROW(c, E + EQ_TORSO_ROT) = QROW(c, L::BQ * 0 + po + 1);
  const V3 velocity = vdiv(ref - ldv(c, h.q_ep), h.env_dt);  // ant.py:226, humanoid.py:252
  const float forward = h.env_kind == BX_ENV_ANT ? velocity.x : h.env_wf * velocity.x;
  float healthy = p0.z < h.env_zmin ? 0.f : 1.f;             // ant.py:229-231
  healthy = p0.z > h.env_zmax ? 0.f : healthy;
  const bool term = (h.env_flags & 2) != 0;
  const float hr = term ? h.env_hr : h.env_hr * healthy;
  for (int i = 0; i < h.act_size; i++) {
    const float a = act_at(c, h, i);
    ctrl += a * a;
  }
  ctrl *= h.env_wc;
  const float dist = bx_sqrt(ref.x * ref.x + ref.y * ref.y + ref.z * ref.z);
  const float done = term ? 1.f - healthy : 0.f;
  if (h.env_kind == BX_ENV_ANT) {
    const float reward = forward + hr - ctrl - contact;        // ant.py:240
    QROW(c, E + EQ_S0) = make_float4(reward, done, forward, hr);
    QROW(c, E + EQ_S1) = make_float4(-ctrl, -contact, ref.x, ref.y);
    QROW(c, E + EQ_S2) = make_float4(dist, velocity.x, velocity.y, forward);
  } else {
    const float reward = forward + hr - ctrl;                  // humanoid.py:265
    QROW(c, E + EQ_S0) = make_float4(reward, done, forward, forward);
    QROW(c, E + EQ_S1) = make_float4(-ctrl, hr, ref.x, ref.y);
    QROW(c, E + EQ_S2) = make_float4(dist, velocity.x, velocity.y, 0.f);
  }
}

// ---- optional in-kernel policy (include/brax_b200.h bx_policy): actions = policy(obs) -----------------
// The reference's acting loop (training/acting.py:53-74) evaluates the policy MLP (training/networks.py:83-102)
// on the normalised obs and feeds NormalTanhDistribution's mode / sample (training/distribution.py:131-158) to
// env.step.  Here: lane = env as everywhere; the warps split the units of a layer in chunks of four (four
// independent accumulators per lane).  A warp stages its 32 rows x 4 units of the weight matrix in its own
// shared-memory slice (the next 32 rows are already in flight in registers), so the inner loop is one broadcast
// LDS.128 of weights + one LDS of the activation per four FMAs and no global-memory latency sits on it.  The
// normalised obs and the activations of a layer sit in shared-memory buffers [unit][lane], a CTA barrier
// separates the layers.  The obs of
// the first step comes from obs0, later steps read the rows the previous step's env epilogue left in shared
// memory through the obs map (envs the auto-reset restarted read first_obs).  Writes the actions into the act
// quads (and act_out).  Called by ALL warps; ends with a barrier.
BXS float policy_act(int kind, float x) {
  if (kind == BX_POLICY_RELU) return fmaxf(x, 0.f);
  if (kind == BX_POLICY_TANH) return tanhf(x);
  // swish: x * sigmoid(x), sigmoid without overflow (exp of a non-positive argument only)
  const float t = __expf(-fabsf(x));
  const float r = bx_rcp(1.f + t);
  return x * (x >= 0.f ? r : t * r);
}
template <int MODE>
static __device__ __forceinline__ void policy_forward(const Ctx c, const BxHeader& h, const BxKernelParams& p, float* pbuf,
                                                   const float* smf, int t, int64_t env0, int nenv, bool restarted) {
  const BxPolicy& q = p.pol;
  const int lane = threadIdx.x & 31;
  const bool live = lane < nenv;
  const int64_t e = env0 + lane;
  const int no = q.sizes[0];
  float* slice = pbuf + c.warp * (4 * BX_POLICY_KBLOCK);                  // this warp's weight rows
  float* obsn = pbuf + BX_POLICY_MAX_WARPS * (4 * BX_POLICY_KBLOCK);      // [obs_rows][32]
  float* hid_a = obsn + q.obs_rows * 32;                                  // [hmax][32] x 2
  float* hid_b = hid_a + q.hmax * 32;
  // Weights do not depend on activations: the first 32-row block (and the bias) of a warp's NEXT chunk --
  // also across the layer barrier -- is fetched into registers while the current chunk computes.
  // element i * 32 + lane of a 32-row block = (row i * 8 + lane / 4, unit lane % 4): a warp load covers
  // 8 rows x 16 bytes.
  float wreg[BX_POLICY_KBLOCK / 8], breg[4];
  auto fetch = [&](const float* wl, int k_hi, int nout, int u0, int k0) {
    const int su = u0 + (lane & 3);
#pragma unroll
    for (int i = 0; i < BX_POLICY_KBLOCK / 8; i++) {
      const int row = k0 + i * 8 + (lane >> 2);
      wreg[i] = (row < k_hi && su < nout) ? __ldg(wl + static_cast<size_t>(row) * nout + su) : 0.f;
    }
  };
  auto fetch_bias = [&](const float* bias, int nout, int u0) {
#pragma unroll
    for (int j = 0; j < 4; j++) breg[j] = u0 + j < nout ? __ldg(bias + u0 + j) : 0.f;
  };
  PTM0(c);
  const int orows = q.obs_rows;
  bool have = 4 * c.warp < q.sizes[1];
  if (have) {
    fetch(q.params, no < orows ? no : orows, q.sizes[1], 4 * c.warp, 0);
    fetch_bias(q.params + static_cast<size_t>(no) * q.sizes[1], q.sizes[1], 4 * c.warp);
  }
  const int na = h.act_size;
  float eps0 = 0.f;   // the noise of this warp's first action, in flight during the whole forward pass
  if (q.noise != nullptr && c.warp < na && live) eps0 = q.noise[(static_cast<int64_t>(t) * p.num_envs + e) * na + c.warp];

  // Rows [k_lo, k_hi) of layer l's input for all of this warp's unit chunks: acc = (first range ? bias :
  // the partial sums left in xout) + sum_k W[k][u] x[k]; the last range applies the activation.
  // xin is indexed by the absolute row.
  auto run_range = [&](int l, const float* w, int nin, int nout, int k_lo, int k_hi, const float* xin, float* xout) {
    const float* bias = w + static_cast<size_t>(nin) * nout;
    const bool first = k_lo == 0, last = k_hi == nin;
    for (int u0 = 4 * c.warp; u0 < nout; u0 += 4 * c.nwarp) {
      if (!have) {
        fetch(w, k_hi, nout, u0, k_lo);
        if (first) fetch_bias(bias, nout, u0);
      }
      have = false;
      // two packed accumulators: FFMA2 (fma.rn.f32x2, sm_100) does two units per issue slot, each
      // component rounded like a scalar fmaf
      float2 acc01, acc23;
      if (first) {
        acc01 = make_float2(breg[0], breg[1]);
        acc23 = make_float2(breg[2], breg[3]);
      } else {
        acc01 = make_float2(xout[u0 * 32 + lane], u0 + 1 < nout ? xout[(u0 + 1) * 32 + lane] : 0.f);
        acc23 = make_float2(u0 + 2 < nout ? xout[(u0 + 2) * 32 + lane] : 0.f, u0 + 3 < nout ? xout[(u0 + 3) * 32 + lane] : 0.f);
      }
      for (int k0 = k_lo; k0 < k_hi; k0 += BX_POLICY_KBLOCK) {
        __syncwarp();
#pragma unroll
        for (int i = 0; i < BX_POLICY_KBLOCK / 8; i++) slice[i * 32 + lane] = wreg[i];
        __syncwarp();
        if (k0 + BX_POLICY_KBLOCK < k_hi) {
          fetch(w, k_hi, nout, u0, k0 + BX_POLICY_KBLOCK);
        } else if (u0 + 4 * c.nwarp < nout) {
          fetch(w, k_hi, nout, u0 + 4 * c.nwarp, k_lo);
          if (first) fetch_bias(bias, nout, u0 + 4 * c.nwarp);
          have = true;
        } else if (last && l + 1 < q.num_layers && 4 * c.warp < q.sizes[l + 2]) {
          fetch(bias + nout, nout, q.sizes[l + 2], 4 * c.warp, 0);
          fetch_bias(bias + nout + static_cast<size_t>(nout) * q.sizes[l + 2], q.sizes[l + 2], 4 * c.warp);
          have = true;
        }
        const int nk = k_hi - k0 < BX_POLICY_KBLOCK ? k_hi - k0 : BX_POLICY_KBLOCK;
        const float* xk = xin + k0 * 32 + lane;
#pragma unroll 4
        for (int kk = 0; kk < nk; kk++) {
          const float4 wv = *reinterpret_cast<const float4*>(slice + 4 * kk);
          const float x = xk[kk * 32];
          const float2 xx = make_float2(x, x);
          acc01 = __ffma2_rn(make_float2(wv.x, wv.y), xx, acc01);
          acc23 = __ffma2_rn(make_float2(wv.z, wv.w), xx, acc23);
        }
      }
      const float acc[4] = {acc01.x, acc01.y, acc23.x, acc23.y};
      const bool act = last && l + 1 < q.num_layers;
#pragma unroll
      for (int j = 0; j < 4; j++)
        if (u0 + j < nout) xout[(u0 + j) * 32 + lane] = act ? policy_act(q.activation, acc[j]) : acc[j];
    }
  };

  // layer 0 over the normalised obs, `obs_rows` rows at a time (all of it when it fits beside the tile; the
  // host shrinks the chunk when that is what lets two tiles share an SM)
  {
    const BxSeg sg = h.seg_out[BX_SEG_ENV];
    const int32_t* omap = c.ti + sg.map;
    const float* obs_g = t == 0 ? p.obs0 : (restarted ? p.first_obs : nullptr);   // global source of this lane's obs
    for (int r0 = 0; r0 < no; r0 += orows) {
      const int r1 = r0 + orows < no ? r0 + orows : no;
      if (r0 > 0) __syncthreads();   // the previous chunk has been consumed
#pragma unroll 4
      for (int k = r0 + c.warp; k < r1; k += c.nwarp) {
        float x;
        if (obs_g != nullptr) {
          x = live ? obs_g[e * no + k] : 0.f;
        } else {
          const int m = omap[k];
          x = m >= 0 ? smf[m + 4 * lane] : sg.fill;
        }
        if (q.obs_mean != nullptr) x -= __ldg(q.obs_mean + k);
        if (q.obs_std != nullptr) x = bx_div(x, __ldg(q.obs_std + k));
        obsn[(k - r0) * 32 + lane] = x;
      }
      __syncthreads();
      if (r0 == 0) PTM(c, 16);
      run_range(0, q.params, no, q.sizes[1], r0, r1, obsn - r0 * 32, hid_a);
    }
  }
  PTM(c, 32);
  __syncthreads();
  PTM(c, 17);
  const float* w = q.params + (static_cast<size_t>(no) + 1) * q.sizes[1];
  const float* xin = hid_a;
  for (int l = 1; l < q.num_layers; l++) {
    const int nin = q.sizes[l], nout = q.sizes[l + 1];
    float* xout = (l & 1) ? hid_b : hid_a;
    run_range(l, w, nin, nout, 0, nin, xin, xout);
    w += (static_cast<size_t>(nin) + 1) * nout;
    xin = xout;
    PTM(c, 32 + l);
    __syncthreads();
    PTM(c, 17 + l);
  }
  // NormalTanhDistribution: loc = out[:A], scale = softplus(out[A:]) + min_std; mode / sample, then tanh
  const float* out = xin;
  for (int a = c.warp; a < na; a += c.nwarp) {
    float loc = out[a * 32 + lane];
    if (q.noise != nullptr) {
      const float raw = out[(na + a) * 32 + lane];
      const float scale = (raw > 20.f ? raw : log1pf(__expf(raw))) + q.min_std;   // softplus
      const float eps = a == c.warp ? eps0 : (live ? q.noise[(static_cast<int64_t>(t) * p.num_envs + e) * na + a] : 0.f);
      loc = fmaf(scale, eps, loc);
    }
    const float act = tanhf(loc);
    *act_ptr(c, h, a) = act;
    if (p.act_out != nullptr && live) p.act_out[(static_cast<int64_t>(t) * p.num_envs + e) * na + a] = act;
  }
  __syncthreads();
  PTM(c, 23);
}

// ============================== the kernel ====================================
// The whole step of one tile of 32 environments (called by bx_step_kernel below).
template <int MODE>
static __device__ __forceinline__ void run(const BxKernelParams& p) {
  typedef Lay<MODE> L;
  const BxHeader& h = p.h;
  float4* sq = bx_smem;
  float4* tf = sq + h.n_quads * QS;
  int32_t* ti = reinterpret_cast<int32_t*>(tf + h.n_fquads);
  float* smf = reinterpret_cast<float*>(sq);
  const int64_t env0 = static_cast<int64_t>(blockIdx.x) * BX_ENVS_PER_CTA;
  const int nenv = static_cast<int>((p.num_envs - env0) < BX_ENVS_PER_CTA ? (p.num_envs - env0) : BX_ENVS_PER_CTA);

  PT_INIT();
  pdl_trigger();
  if (p.stagger_ns) {
    // All tiles of a small batch are resident at once and would move in lock step:
    // every tile reading over PCIe, then every tile writing.  Spreading the starts
    // over the transfer time keeps both directions of the link busy together.
    if (threadIdx.x == 0) {
      const unsigned long long wait =
          static_cast<unsigned long long>(blockIdx.x) * p.stagger_ns / gridDim.x;
      unsigned long long t0, t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
      for (;;) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 >= wait) break;
        __nanosleep(static_cast<unsigned>(wait - (t - t0) < 1000ull ? wait - (t - t0) : 1000ull));
      }
    }
    __syncthreads();
  }
  // ---- stage in: ONE wave of bulk async copies (tables + the five input spans of
  // this tile), one mbarrier, then the transposition staging -> state quads.
  uint64_t* mbar = reinterpret_cast<uint64_t*>(ti + h.n_ints);
  float* stage = smf + h.q_js * QS * 4;  // JS.. regions are dead outside the substep loop
  const bool full = nenv == BX_ENVS_PER_CTA;
  // env builds with an in-kernel policy: the actions are computed in the launch, `act` is not read
  const bool policy_on = (BX_FEAT & BXF_ENV) != 0 && p.pol.num_layers > 0;
  const int n_in = policy_on ? BX_SEG_IN - 1 : BX_SEG_IN;   // (act is the last input segment)
  bool in_bulk = full;
#pragma unroll
  for (int s = 0; s < BX_SEG_IN; s++) in_bulk = in_bulk && (s >= n_in || span_bulk_ok(h.seg_in[s], p.in[s], env0));
  uint32_t phase = 0;
  if (threadIdx.x == 0) mbar_init(mbar, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t bytes = h.n_fquads * 16 + h.n_ints * 4;
    if (in_bulk)
      for (int s = 0; s < n_in; s++) bytes += 4 * BX_ENVS_PER_CTA * h.seg_in[s].nper;
    mbar_expect_tx(mbar, bytes);
    bulk_g2s(tf, p.tab_f, h.n_fquads * 16, mbar);  // static tables: not produced by a kernel
    bulk_g2s(ti, p.tab_i, h.n_ints * 4, mbar);
    pdl_wait();  // the inputs may be the previous launch's outputs
    if (in_bulk) {
      int off = 0;
      for (int s = 0; s < n_in; s++) {
        const int n = BX_ENVS_PER_CTA * h.seg_in[s].nper;
        if (n) bulk_g2s(stage + off, p.in[s] + env0 * h.seg_in[s].nper, 4 * n, mbar);
        off += n;
      }
    }
  }
  if (threadIdx.x != 0) pdl_wait();  // every thread that will read or write global memory
  if (h.fr_off && threadIdx.x >= blockDim.x - 32) {
    // per-env collider friction: a pair's friction is friction_a * friction_b = friction[e]^2 (float32
    // product, like the descriptor's); read by the contact rows after the barriers that follow
    const int l = threadIdx.x & 31;
    const float f = l < nenv ? p.friction[env0 + l] : 1.f;
    smf[h.fr_off + l] = f * f;
  }
  if ((BX_FEAT & BXF_JOINT_OFFSET) != 0 && h.jo_off) {
    // per-env joint sockets: Joint.parent_offset of env `lane` for every joint, read by the joint phases
    for (int i = threadIdx.x; i < h.nj * BX_ENVS_PER_CTA; i += blockDim.x) {
      const int j = i / BX_ENVS_PER_CTA, l = i - j * BX_ENVS_PER_CTA;
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      if (l < nenv) {
        const float* src = p.joint_offset + ((env0 + l) * h.nj + j) * 3;
        o = make_float4(src[0], src[1], src[2], 0.f);
      }
      bx_smem[h.jo_off + i] = o;
    }
  }
  if (!full) {
    // partial tile: give the idle env slots a benign state (identity rotations)
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = threadIdx.x; i < h.n_quads * QS; i += blockDim.x) sq[i] = z;
    __syncthreads();
    const int nrot = MODE == BX_MODE_PBD ? 2 : 1;
    for (int i = threadIdx.x; i < h.nb * nrot * BX_ENVS_PER_CTA; i += blockDim.x) {
      const int br = i / BX_ENVS_PER_CTA, e = i - br * BX_ENVS_PER_CTA;
      const int b = br / nrot, which = br - b * nrot;
      sq[(L::BQ * b + 2 * which + 1) * QS + e] = make_float4(1.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
  }
  mbar_wait(mbar, phase);
  phase ^= 1;
  PT(0);
  if (in_bulk) {
    for (int s = threadIdx.x >> 5; s < n_in; s += blockDim.x >> 5) {
      int off = 0;
      for (int t = 0; t < s; t++) off += BX_ENVS_PER_CTA * h.seg_in[t].nper;
      unstage_warp(smf, ti, h.seg_in[s], stage + off, threadIdx.x & 31);
    }
  } else {
#pragma unroll 1
    for (int s = 0; s < n_in; s++)
      if (h.seg_in[s].nper)
        io_segment_scalar<true>(smf, ti, h.seg_in[s], const_cast<float*>(p.in[s]), env0, nenv);
  }

  Ctx c;
  c.warp = threadIdx.x >> 5;
  c.nwarp = blockDim.x >> 5;
  c.q = sq + (threadIdx.x & 31);
  c.tf = tf;
  c.ti = ti;
  c.cur = 0;
  {
    // the damping items of phase T are short: when the last actuator round leaves
    // warps idle, those warps take them (up to 4 each) instead of the busy ones
    const int rem = h.na % c.nwarp, n_idle = rem ? c.nwarp - rem : 0;
    const bool on_idle = n_idle > 0 && h.nj <= 4 * n_idle;
    c.damp_j0 = on_idle ? (c.warp >= rem ? c.warp - rem : h.nj) : c.warp;
    c.damp_step = on_idle ? n_idle : c.nwarp;
  }
#ifdef BX_PHASE_TIMING
  c.dbg = p.dbg;
  c.mark = 0;
  c.moff = 0;
#endif

  const bool env_on = (BX_FEAT & BXF_ENV) != 0 && h.env_kind != BX_ENV_NONE && (p.out[BX_SEG_ENV] != nullptr || p.out[BX_SEG_ENV + 1] != nullptr ||
                                                    p.out[BX_SEG_ENV + 2] != nullptr || p.out[BX_SEG_ENV + 3] != nullptr);
  // auto-reset (EpisodeWrapper + AutoResetWrapper around the env epilogue, wrappers.py:105-148):
  // the step counter and the previous done flag of env `lane` live in registers of the last warp
  const bool auto_reset = env_on && p.first[0] != nullptr;
  float ep_steps = 0.f, ep_prev_done = 0.f;
  if (auto_reset && c.warp == c.nwarp - 1 && (threadIdx.x & 31) < nenv) {
    ep_prev_done = p.done_in[env0 + (threadIdx.x & 31)];
    if (p.steps_in != nullptr) ep_steps = p.steps_in[env0 + (threadIdx.x & 31)];
  }
  // observe (env builds): zero substeps -- the epilogue then sees the state passed in and a zero Info
  const int iters = ((BX_FEAT & BXF_ENV) != 0 && p.observe != 0) ? 0 : h.iters;
  // EpisodeWrapper's action_repeat (wrappers.py:106-110): `repeat` env steps share one action, their
  // rewards are summed, and the wrappers' bookkeeping / outputs happen after the last of them
  const int repeat = (auto_reset && p.action_repeat > 1) ? p.action_repeat : 1;
  float ep_reward = 0.f;
  for (int step = 0; step < p.num_steps; step++) {
    const int ot = step / repeat, rep = step - ot * repeat;
    __syncthreads();
    if (policy_on) {
      // actions of this step from the policy: obs of the previous step (obs0 for the first; first_obs for
      // envs the auto-reset restarted at the end of the previous step)
      // (the policy's buffers start at the 16-byte boundary behind the mbarriers)
      float* pbuf = reinterpret_cast<float*>(bx_smem + h.n_quads * QS + h.n_fquads + ((h.n_ints * 4 + 16 + BX_FRICTION_SMEM + 15) >> 4) +
                                             (h.jo_off ? h.nj * BX_ENVS_PER_CTA : 0));
      const bool restarted = step > 0 && p.first[0] != nullptr && (threadIdx.x & 31) < nenv &&
                             sq[(h.q_env + EQ_S0) * QS + (threadIdx.x & 31)].y > 0.f;
      policy_forward<MODE>(c, h, p, pbuf, smf, step, env0, nenv, restarted);
    } else if (step > 0 && rep == 0) {  // rollout: the actions of step t (step 0 came with the state)
      const float* act = p.in[4] + static_cast<int64_t>(ot) * p.num_envs * h.act_size;
      const BxSeg sg = h.seg_in[4];
      if (full && span_bulk_ok(sg, act, env0)) {
        if (threadIdx.x == 0) {
          fence_async_smem();
          mbar_expect_tx(mbar, 4 * BX_ENVS_PER_CTA * sg.nper);
          bulk_g2s(stage, act + env0 * sg.nper, 4 * BX_ENVS_PER_CTA * sg.nper, mbar);
        }
        mbar_wait(mbar, phase);
        phase ^= 1;
        if (threadIdx.x < 32) unstage_warp(smf, ti, sg, stage, threadIdx.x);
      } else {
        io_segment_scalar<true>(smf, ti, sg, const_cast<float*>(act), env0, nenv);
      }
      __syncthreads();
    }
    // zero_info (system.py:76-84): Info accumulators start at 0 every step
    {
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int i = threadIdx.x; i < h.n_zero * QS; i += blockDim.x) sq[h.q_ic * QS + i] = z;
      if (iters == 0)  // contact_penetration = -1 (system.py:84)
        for (int i = threadIdx.x; i < (h.q_pre - h.q_ct) * QS; i += blockDim.x)
          sq[h.q_ct * QS + i] = i < h.nk * QS ? make_float4(0.f, 0.f, 0.f, -1.f) : z;
      // (observe: grasp's prologue -- action map and palm move, grasp.py:75-85 -- belongs to step)
      if (env_on && !(p.observe != 0 && h.env_kind == BX_ENV_GRASP)) phase_env_before<MODE>(c, h);
    }
    __syncthreads();
    PT(1);
    if constexpr (MODE == BX_MODE_PBD) {
      for (int it = 0; it < iters; it++) {
#pragma unroll 1
        for (int half = 0; half < 2; half++) {
          phase_torques<MODE>(c, h, true);
          PTA(2); __syncthreads(); PT(2);
          phase_body_acc_kinetic(c, h, half == 1);
#ifdef BX_EXPERIMENT_TWICE  // same phase again, instruction cache warm: timing only
          __syncthreads();
          c.moff = 16;
          phase_body_acc_kinetic(c, h, half == 1);
          c.moff = 0;
#endif
          c.cur ^= 1;
          PTA(3); __syncthreads(); PT(3);
          phase_joints_pbd(c, h);
          PTA(4); __syncthreads(); PT(4);
          phase_body_joint_update(c, h, half == 0);
          PTA(5); __syncthreads(); PT(5);
#ifndef BX_EXPERIMENT_NO_CONTACTS
          if (half == 1) {
            if (BX_FEAT & BXF_FUSED_CONTACTS) {
              phase_contacts_fused(c, h);
              PTA(6); __syncthreads(); PT(6);
            } else {
              phase_contacts_position(c, h);
              PTA(6); __syncthreads(); PT(6);
              phase_body_contact_update(c, h);
              PTA(7); __syncthreads(); PT(7);
              phase_contacts_velocity(c, h);
              PTA(8); __syncthreads(); PT(8);
              phase_body_velocity_update<MODE>(c, h, 1e-6f);
              PTA(9); __syncthreads(); PT(9);
            }
          }
#endif
        }
      }
    } else {
      for (int it = 0; it < iters; it++) {
        phase_body_kinetic(c, h);
        __syncthreads();
        phase_torques<MODE>(c, h, false);
        phase_joints_spring(c, h);
        __syncthreads();
        phase_body_acc_spring(c, h);
        __syncthreads();
        phase_contacts_spring<MODE>(c, h);
        __syncthreads();
        phase_body_velocity_update<MODE>(c, h, 1e-8f);
        // same warp<->body mapping as phase_body_kinetic: no barrier needed
      }
    }
    if ((BX_FEAT & BXF_ENV) != 0 && p.observe != 0 && h.nk > 0) {
      // observe (Env.reset): info = sys.info(qp) of the state passed in (envs/ant.py:204) -- the
      // contact rows and the spring-style contact impulse; actuator / joint parts are not read by any obs
      __syncthreads();
      phase_contacts_spring<MODE>(c, h);
      __syncthreads();
      phase_info_bodies<MODE>(c, h);
    }
    if (env_on) {
      // Env.step's epilogue of THIS step (every step of a rollout): obs / reward / done /
      // metrics [T,B,...]
      __syncthreads();
      if (h.env_flags & 4) {
        phase_angle_vel<MODE>(c, h);  // writes rows only the stores read: no barrier before phase_env
        phase_env<MODE>(c, h);
      } else {
        phase_env_ego<MODE>(c, h, p.env_rand == nullptr ? nullptr
                            : p.env_rand + 3 * (static_cast<int64_t>(step) * p.num_envs + env0), nenv);
      }
      if (auto_reset) {
        const int lane = threadIdx.x & 31;
        if (c.warp == c.nwarp - 1) {
          // EpisodeWrapper.step / AutoResetWrapper.step bookkeeping of env `lane` (the same warp
          // wrote the scalars row in phase_env*)
          float4 s0 = QROW(c, h.q_env + EQ_S0);
          ep_reward = (rep == 0 ? 0.f : ep_reward) + s0.x;            // jp.sum(rewards) wrappers.py:110-111
          if (rep == repeat - 1) {
            s0.x = ep_reward;
            ep_steps = ep_prev_done > 0.f ? 0.f : ep_steps;           // wrappers.py:133-135
            ep_steps += static_cast<float>(repeat);                   // wrappers.py:112
            const bool over = p.episode_length > 0.f && ep_steps >= p.episode_length;
            const float trunc = over ? 1.f - s0.y : 0.f;              // wrappers.py:117-118
            s0.y = over ? 1.f : s0.y;                                 // wrappers.py:116
            ep_prev_done = s0.y;
            QROW(c, h.q_env + EQ_S0) = s0;
            if (p.truncation != nullptr && lane < nenv)
              p.truncation[static_cast<int64_t>(ot) * p.num_envs + env0 + lane] = trunc;
          }
        }
        if (rep < repeat - 1) continue;  // (whole CTA) the env's next step with the same action
        __syncthreads();
        // qp = where(done, first_qp, qp) (wrappers.py:146): done envs reload their first state
        const bool done = lane < nenv && QROW(c, h.q_env + EQ_S0).y > 0.f;
        if (__any_sync(0xffffffffu, done)) {
          typedef Lay<MODE> L;
          const int po = MODE == BX_MODE_PBD ? 2 * c.cur : 0;
          const int64_t e = env0 + lane;
          for (int b = c.warp; b < h.nb; b += c.nwarp) {
            if (!done) continue;
            const float* fp = p.first[0] + (e * h.nb + b) * 3;
            const float* fr = p.first[1] + (e * h.nb + b) * 4;
            const float* fv = p.first[2] + (e * h.nb + b) * 3;
            const float* fa = p.first[3] + (e * h.nb + b) * 3;
            QROW(c, L::BQ * b + po) = make_float4(fp[0], fp[1], fp[2], 0.f);
            QROW(c, L::BQ * b + po + 1) = make_float4(fr[0], fr[1], fr[2], fr[3]);
            QROW(c, L::BQ * b + L::VEL) = make_float4(fv[0], fv[1], fv[2], 0.f);
            QROW(c, L::BQ * b + L::ANG) = make_float4(fa[0], fa[1], fa[2], 0.f);
          }
        }
      }
      __syncthreads();
      auto env_ptr = [&](int s) -> float* {
        return p.out[s] == nullptr ? nullptr
            : p.out[s] + static_cast<int64_t>(ot) * p.num_envs * h.seg_out[s].nper;
      };
      bool vec = full;
#pragma unroll
      for (int s = BX_SEG_ENV; s < BX_SEG_OUT; s++)
        vec = vec && (p.out[s] == nullptr || span_bulk_ok(h.seg_out[s], env_ptr(s), env0));
      if (vec) {
        // obs is long (87 .. 240 floats per env): the whole CTA shares it; the three short
        // segments go to the last warps
        if (p.out[BX_SEG_ENV] != nullptr)
          store_warp(smf, ti, h.seg_out[BX_SEG_ENV], env_ptr(BX_SEG_ENV) + env0 * h.seg_out[BX_SEG_ENV].nper,
                     threadIdx.x, blockDim.x);
        for (int s = BX_SEG_ENV + 1 + (c.nwarp - 1 - c.warp); s < BX_SEG_OUT; s += c.nwarp)
          if (p.out[s] != nullptr)
            store_warp(smf, ti, h.seg_out[s], env_ptr(s) + env0 * h.seg_out[s].nper, threadIdx.x & 31);
      } else {
#pragma unroll 1
        for (int s = BX_SEG_ENV; s < BX_SEG_OUT; s++)
          if (p.out[s] != nullptr)
            io_segment_scalar<false>(smf, ti, h.seg_out[s], env_ptr(s), env0, nenv);
      }
      if (auto_reset && p.out[BX_SEG_ENV] != nullptr) {
        // obs = where(done, first_obs, obs) (wrappers.py:147): after this tile's obs store; the
        // barrier also tells whether any env of the tile is done at all
        const int ln = threadIdx.x & 31;
        const int any_done = __syncthreads_or(ln < nenv && sq[(h.q_env + EQ_S0) * QS + ln].y > 0.f);
        if (any_done) {
          // (one flat pass over the tile's [nenv, O] span: the copies of different envs are
          // independent loads in flight together, not one load -> store round trip per done env)
          const int no = h.seg_out[BX_SEG_ENV].nper;
          float* __restrict__ obs = env_ptr(BX_SEG_ENV) + env0 * no;
          const float* __restrict__ fobs = p.first_obs + env0 * no;
          const int total = nenv * no;
#pragma unroll 4
          for (int i = threadIdx.x; i < total; i += blockDim.x) {
            const int e = i / no;
            if (sq[(h.q_env + EQ_S0) * QS + e].y > 0.f) obs[i] = __ldg(fobs + i);
          }
        }
      }
    }
  }
  if (auto_reset && p.steps_out != nullptr && c.warp == c.nwarp - 1 && (threadIdx.x & 31) < nenv)
    p.steps_out[env0 + (threadIdx.x & 31)] = ep_steps;
  __syncthreads();
  if (!env_on && (p.out[13] != nullptr || p.out[14] != nullptr)) {
    phase_angle_vel<MODE>(c, h);
    __syncthreads();
  }

  PT(10);
  // ---- out: each warp transposes whole segments straight to HBM with coalesced
  // 128-bit stores.  The pos/rot maps point at buffer 0; buffer `cur` holds the
  // result.
  const int cur_shift = (MODE == BX_MODE_PBD && c.cur) ? 2 * QS * 4 : 0;
  bool out_vec = full;
#pragma unroll
  for (int s = 0; s < BX_SEG_OUT_STEP; s++)
    out_vec = out_vec && (p.out[s] == nullptr || span_bulk_ok(h.seg_out[s], p.out[s], env0));
  if (out_vec) {
    for (int s = c.warp; s < BX_SEG_OUT_STEP; s += c.nwarp)
      if (p.out[s] != nullptr && h.seg_out[s].nper > 0)
        store_warp(smf + (s < 2 ? cur_shift : 0), ti, h.seg_out[s],
                   p.out[s] + env0 * h.seg_out[s].nper, threadIdx.x & 31);
  } else {
#pragma unroll 1
    for (int s = 0; s < BX_SEG_OUT_STEP; s++)
      if (p.out[s] != nullptr && h.seg_out[s].nper > 0)
        io_segment_scalar<false>(smf + (s < 2 ? cur_shift : 0), ti, h.seg_out[s], p.out[s], env0, nenv);
  }
  PT(11);
  PT_DONE();
}
};  // struct Kern

// blockDim.x = 32 * W (W warps share one tile of 32 environments).  MAXT / MINB
// are the launch bounds: the register budget is what keeps two tiles per SM
// resident (256 threads: 128 regs, 320: 102 regs).
template <int MODE, int FEAT, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB)
bx_step_kernel(const __grid_constant__ BxKernelParams p) {
  Kern<FEAT>::template run<MODE>(p);
}
