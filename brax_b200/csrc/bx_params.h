// bx_params.h -- kernel parameter block, feature masks and the launcher
// interface between bx_api.cu (host API) and the kernel translation units.
#ifndef BX_PARAMS_H_
#define BX_PARAMS_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include "bx_tables.h"

// bx_policy by value (include/brax_b200.h): the in-kernel MLP policy of a rollout
// reset joint j (depth order of bx_reset_desc.joints) -> its index in bx_system_desc.joints: how the reset kernel
// finds a joint's socket in a per-env joint_offset[B,J,3] array (bx_default_qp_offsets)
struct BxResetMap {
  int32_t num_joints;
  int32_t step_joint[64];
};
#define BX_FRICTION_SMEM 128   // bytes behind the mbarriers: friction[e]^2 of the tile's 32 envs
#define BX_POLICY_KBLOCK 32     // weight rows a warp stages per block (4 units wide)
#define BX_POLICY_MAX_WARPS 16  // slices are laid out for the widest launch
// shared memory of the in-kernel policy beside the tile: weight slices, normalised obs, two activation buffers
#define BX_POLICY_SMEM(obs_rows, hmax) \
  (16 + sizeof(float) * (BX_POLICY_MAX_WARPS * 4 * BX_POLICY_KBLOCK + 32 * ((obs_rows) + 2 * (hmax))))
struct BxPolicy {
  int32_t num_layers;  // 0 = no policy
  int32_t sizes[9];
  int32_t activation;
  int32_t hmax;        // widest layer output: the two activation buffers hold hmax x 32 floats each
  int32_t obs_rows;    // rows of the normalised obs staged at a time (= sizes[0] when all of it fits)
  float min_std;
  const float* params;
  const float* obs_mean;
  const float* obs_std;
  const float* noise;
};

struct BxKernelParams {
  BxHeader h;
  const int32_t* tab_i;
  const float4* tab_f;
  const float* in[BX_SEG_IN];  // pos rot vel ang act
  float* out[BX_SEG_OUT];      // order of bx_step_io's outputs
  const float* env_rand;       // fetch / grasp: [T,B,3] uniforms of the target teleport, or NULL
  // auto-reset (EpisodeWrapper + AutoResetWrapper, wrappers.py:105-148); first[0] NULL = off
  const float* first[4];       // first_qp: pos rot vel ang
  const float* first_obs;
  const float* done_in;
  const float* steps_in;
  float* steps_out;
  float* truncation;
  float episode_length;
  int observe;                 // env builds: skip the substeps (obs of the state passed in)
  int action_repeat;           // EpisodeWrapper.action_repeat (auto-reset launches), >= 1
  BxPolicy pol;                // in-kernel policy (env builds), pol.num_layers == 0: none
  const float* obs0;           // [B,O] obs of the state passed in (policy launches)
  float* act_out;              // [T,B,A] actions the policy chose, or NULL
  const float* friction;       // [B] per-env collider friction (h.fr_off != 0), or NULL
  const float* joint_offset;   // [B,J,3] per-env Joint parent_offset (h.jo_off != 0), or NULL
  int64_t num_envs;
  int32_t num_steps;  // >1: rollout, act is [T,B,A]
  // host path only (bx_step_host on mapped host memory): tile i starts i/grid of
  // this many nanoseconds late, so the tiles' PCIe reads and writes overlap (else 0)
  uint32_t stagger_ns;
  long long* dbg;     // BX_PHASE_TIMING builds only: cycles per phase of CTA 0
};


// Feature mask of a kernel variant: code for features a system does not use is
// compiled out (smaller instruction footprint of the substep loop).
enum {
  BXF_REVOLUTE = 1, BXF_SPHERICAL = 2, BXF_UNIVERSAL = 4,
  BXF_ONE_WAY = 8, BXF_TWO_WAY = 16,
  BXF_CAPSULE_PLANE = 32, BXF_CAPSULE_CAPSULE = 64, BXF_BOX_PLANE = 128,
  BXF_TORQUE = 256, BXF_ANGLE = 512, BXF_FORCES = 1024,
  BXF_FUSED_CONTACTS = 2048,  // pbd, one-way colliders, <= 2 rows per body
  BXF_ENV = 4096,             // env epilogue compiled in (systems created with desc.env)
  BXF_CAPSULE_BOX = 8192,     // capsule against a triangulated box (capsule_mesh, colliders.py:822-849)
  BXF_JOINT_OFFSET = 16384,   // per-env joint sockets (bx_step_io.joint_offset): all-contact-features builds of their own
  BXF_ALL = 2047
};
#define BX_CC_ENV 8  // or-ed into the contact class passed to bx_launch_*: env-epilogue build
// or-ed in as well: launch with programmatic stream serialization (programmatic dependent
// launch): the CTAs of this launch may become resident, load the static tables and park at
// griddepcontrol.wait while the previous kernel of the stream drains
#define BX_CC_PDL 16
// or-ed in as well: the launch carries per-env joint sockets -> the build with BXF_JOINT_OFFSET of the system's
// joint class (contact class MESH = every contact feature; only in the translation units that hold the MESH class)
#define BX_CC_JO 32
// Kernel variants are instantiated per (dynamics mode, joint class, contact
// class, launch bound).  Classes are the feature sets of the reference's envs (BX_CC_MESH: the
// round-2 capsule-box collider on top of everything):
// all-revolute (ant, fetch), all-spherical (humanoid, grasp: joints.py:418-474
// "sphericalize"), anything; plane contacts only (one-way capsule/box vs plane)
// or everything.
// The joint class also fixes the actuator features: the revolute and spherical classes are torque-actuated without
// forces (27 of the reference's 31 env configs), spherical + angle actuators is grasp's, anything else -- angle
// actuators on revolute joints, thrusters / twisters, mixed joint types, spring-mode universal joints -- runs ALL.
// (Measured: compiling the unused angle-actuator / force code out of the torque-only builds is worth 0.3 % on ant,
// 1.5 % on fetch and 2.2 % on humanoid: less code in the substep loop.)
enum { BX_JC_REVOLUTE = 0, BX_JC_SPHERICAL = 1, BX_JC_ALL = 2, BX_JC_SPHERICAL_ANGLE = 3, BX_NUM_JC = 4 };
enum { BX_CC_PLANE_FUSED = 0, BX_CC_PLANE = 1, BX_CC_ALL = 2, BX_CC_MESH = 3, BX_NUM_CC = 4 };
#define BXF_COMMON 0
#define BXF_JC0 (BXF_REVOLUTE | BXF_TORQUE)
#define BXF_JC1 (BXF_SPHERICAL | BXF_TORQUE)
#define BXF_JC2 (BXF_REVOLUTE | BXF_SPHERICAL | BXF_UNIVERSAL | BXF_TORQUE | BXF_ANGLE | BXF_FORCES)
#define BXF_JC3 (BXF_SPHERICAL | BXF_ANGLE)
#define BXF_CC0 (BXF_ONE_WAY | BXF_CAPSULE_PLANE | BXF_BOX_PLANE)
#define BXF_CC1 (BXF_ONE_WAY | BXF_TWO_WAY | BXF_CAPSULE_PLANE | BXF_CAPSULE_CAPSULE | BXF_BOX_PLANE)
#define BXF_CC2 (BXF_CC1 | BXF_CAPSULE_BOX)  // ... and capsule against a triangulated box

// One object per launch bound, dynamics mode and env-epilogue flag (bx_kernels.cu compiled with -DBX_TU_MAXT=..
// -DBX_TU_MINB=.. -DBX_TU_MODE=0|1 -DBX_TU_ENV=0|1) exports a launcher and a shared-memory setter;
// bx_launch_<bound> picks the object.
#define BX_LAUNCHER_PROTO(sfx)                                                                          \
  cudaError_t bx_launch_##sfx(int mode, int jc, int cc, unsigned grid, int threads, size_t smem,        \
                              cudaStream_t stream, const BxKernelParams& p);                            \
  cudaError_t bx_set_smem_##sfx(int bytes);
#define BX_LAUNCHER_DECL(maxt)                                                                          \
  BX_LAUNCHER_PROTO(maxt##_m0e0) BX_LAUNCHER_PROTO(maxt##_m0e1)                                         \
  BX_LAUNCHER_PROTO(maxt##_m1e0) BX_LAUNCHER_PROTO(maxt##_m1e1)                                         \
  static inline cudaError_t bx_launch_##maxt(int mode, int jc, int cc, unsigned grid, int threads,      \
                                             size_t smem, cudaStream_t stream, const BxKernelParams& p) { \
    if (cc & BX_CC_ENV)                                                                                 \
      return mode == 0 ? bx_launch_##maxt##_m0e1(mode, jc, cc, grid, threads, smem, stream, p)          \
                       : bx_launch_##maxt##_m1e1(mode, jc, cc, grid, threads, smem, stream, p);         \
    return mode == 0 ? bx_launch_##maxt##_m0e0(mode, jc, cc, grid, threads, smem, stream, p)            \
                     : bx_launch_##maxt##_m1e0(mode, jc, cc, grid, threads, smem, stream, p);           \
  }                                                                                                     \
  static inline cudaError_t bx_set_smem_##maxt(int bytes) {                                             \
    cudaError_t e = bx_set_smem_##maxt##_m0e0(bytes);                                                   \
    if (e == cudaSuccess) e = bx_set_smem_##maxt##_m0e1(bytes);                                         \
    if (e == cudaSuccess) e = bx_set_smem_##maxt##_m1e0(bytes);                                         \
    return e != cudaSuccess ? e : bx_set_smem_##maxt##_m1e1(bytes);                                     \
  }
BX_LAUNCHER_DECL(256)
BX_LAUNCHER_DECL(320)
BX_LAUNCHER_DECL(320x3)  // (320, 3): 64 registers, three tiles per SM for big batches of small systems
BX_LAUNCHER_DECL(384)
BX_LAUNCHER_DECL(512)

#endif  // BX_PARAMS_H_
