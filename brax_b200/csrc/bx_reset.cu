// bx_reset.cu -- System.default_qp(joint_angle, joint_velocity) (brax/physics/system.py:112-242)
// for a batch of environments: forward kinematics down the joint tree and the lowering of free
// trees onto the xy plane.  One warp per 32 envs, lane = env; the tile's poses live in shared
// memory as rows [10 per body: pos 3, rot 4, ang 3][32 lanes] and leave through coalesced
// stores of the tile's contiguous [32, N, w] output spans.  The static tables (11 KB) are copied
// into shared memory first: the joint / probe loops are one warp's serial chain of table reads.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/brax_b200.h"
#include "bx_math.cuh"
#include "bx_params.h"

#define ROW(r) sm[(r) * 32 + lane]

__global__ void __launch_bounds__(32)
bx_reset_kernel(const bx_reset_desc* __restrict__ gt, int nb, int64_t num_envs,
                const float* __restrict__ joint_angle, const float* __restrict__ joint_velocity,
                const float* __restrict__ joint_offset, BxResetMap map,
                float* __restrict__ pos_o, float* __restrict__ rot_o, float* __restrict__ vel_o,
                float* __restrict__ ang_o) {
  extern __shared__ float smem_all[];
  const int lane = threadIdx.x;
  constexpr int TAB_WORDS = (sizeof(bx_reset_desc) + 15) / 16 * 4;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(gt);
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem_all);
    for (int i = lane; i < static_cast<int>(sizeof(bx_reset_desc) / 4); i += 32) dst[i] = src[i];
  }
  __syncwarp();
  const bx_reset_desc* t = reinterpret_cast<const bx_reset_desc*>(smem_all);
  float* sm = smem_all + TAB_WORDS;
  const int64_t env0 = static_cast<int64_t>(blockIdx.x) * 32;
  const int nenv = static_cast<int>(num_envs - env0 < 32 ? num_envs - env0 : 32);
  const int64_t e = env0 + (lane < nenv ? lane : 0);
  const int nd = t->num_dof;
  for (int b = 0; b < nb; b++) {  // system.py:118-131
    for (int x = 0; x < 3; x++) ROW(b * 10 + x) = t->base_pos[b][x];
    for (int x = 0; x < 4; x++) ROW(b * 10 + 3 + x) = t->base_rot[b][x];
    for (int x = 0; x < 3; x++) ROW(b * 10 + 7 + x) = t->base_ang[b][x];
  }
  auto body_pos = [&](int b) { return v3(ROW(b * 10), ROW(b * 10 + 1), ROW(b * 10 + 2)); };
  auto body_rot = [&](int b) { return q4(ROW(b * 10 + 3), ROW(b * 10 + 4), ROW(b * 10 + 5), ROW(b * 10 + 6)); };
  for (int j = 0; j < t->num_joints; j++) {
    const bx_reset_joint& jt = t->joints[j];
    float a[3], w[3];
    for (int d = 0; d < 3; d++) {  // to_3dof, system.py:156-168
      const int k = jt.take[d];
      a[d] = k < nd ? joint_angle[e * nd + k] : 0.f;
      w[d] = k < nd ? joint_velocity[e * nd + k] : 0.f;
    }
    const Q4 jr = q4(jt.rot[0], jt.rot[1], jt.rot[2], jt.rot[3]);
    V3 axes[3] = {rotate(v3(1.f, 0.f, 0.f), jr), rotate(v3(0.f, 1.f, 0.f), jr), rotate(v3(0.f, 0.f, 1.f), jr)};
    // ang = dot(axes.T, vels)  (system.py:183)
    const V3 la = v3(axes[0].x * w[0] + axes[1].x * w[1] + axes[2].x * w[2],
                     axes[0].y * w[0] + axes[1].y * w[1] + axes[2].y * w[2],
                     axes[0].z * w[0] + axes[1].z * w[1] + axes[2].z * w[2]);
    Q4 lr = q4(jt.ref[0], jt.ref[1], jt.ref[2], jt.ref[3]);
    for (int d = 0; d < 3; d++)  // intrinsic euler rotations (system.py:185-189)
      lr = quat_mul(quat_rot_axis_any(rotate(axes[d], lr), a[d]), lr);
    const Q4 rp = body_rot(jt.body_p);
    const V3 pp = body_pos(jt.body_p);
    const Q4 wr = quat_mul(rp, lr);                                                     // :204
    // this env's own socket when per-env joint offsets are given ([B,J,3] in the order of bx_system_desc.joints)
    V3 op = v3(jt.off_p[0], jt.off_p[1], jt.off_p[2]);
    if (joint_offset != nullptr) {
      const float* o = joint_offset + (e * map.num_joints + map.step_joint[j]) * 3;
      op = v3(o[0], o[1], o[2]);
    }
    const V3 lp = op - rotate(v3(jt.off_c[0], jt.off_c[1], jt.off_c[2]), lr);
    const V3 wp = pp + rotate(lp, rp);                                                  // :206
    const V3 wa = rotate(la, rp);                                                       // :207
    const int c = jt.body_c * 10;
    ROW(c) = wp.x; ROW(c + 1) = wp.y; ROW(c + 2) = wp.z;
    ROW(c + 3) = wr.w; ROW(c + 4) = wr.x; ROW(c + 5) = wr.y; ROW(c + 6) = wr.z;
    ROW(c + 7) = wa.x; ROW(c + 8) = wa.y; ROW(c + 9) = wa.z;
  }
  for (int g = 0; g < t->num_trees; g++) {  // system.py:211-240, bodies.py:62-98
    float min_z = t->tree_zero[g] ? 0.f : __int_as_float(0x7f800000);
    for (int i = 0; i < t->num_probes; i++) {
      const bx_reset_probe& pr = t->probes[i];
      if (pr.tree != g) continue;
      const V3 pt = rotate(v3(pr.point[0], pr.point[1], pr.point[2]), body_rot(pr.body));
      min_z = fminf(min_z, ROW(pr.body * 10 + 2) + pt.z - pr.radius);
    }
    for (int b = 0; b < nb; b++)
      if (t->body_tree[b] == g) ROW(b * 10 + 2) = ROW(b * 10 + 2) - min_z;
  }
  __syncwarp();
  // coalesced stores of the tile's contiguous output spans
  const int off[4] = {0, 3, -1, 7}, wid[4] = {3, 4, 3, 3};
  float* outs[4] = {pos_o, rot_o, vel_o, ang_o};
  for (int s = 0; s < 4; s++) {
    const int per = nb * wid[s];
    float* o = outs[s] + env0 * per;
    for (int i = lane; i < nenv * per; i += 32) {
      const int el = i / per, k = i - el * per, b = k / wid[s], x = k - b * wid[s];
      o[i] = off[s] < 0 ? t->base_vel[b][x] : sm[(b * 10 + off[s] + x) * 32 + el];
    }
  }
}

cudaError_t bx_reset_launch(const bx_reset_desc* d_tables, int nb, int64_t num_envs, const float* joint_angle,
                            const float* joint_velocity, const float* joint_offset, const BxResetMap& map,
                            float* pos, float* rot, float* vel, float* ang, cudaStream_t stream) {
  const size_t smem = (sizeof(bx_reset_desc) + 15) / 16 * 16 + static_cast<size_t>(nb) * 10 * 32 * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(bx_reset_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
    if (e != cudaSuccess) return e;
  }
  const unsigned grid = static_cast<unsigned>((num_envs + 31) / 32);
  bx_reset_kernel<<<grid, 32, smem, stream>>>(d_tables, nb, num_envs, joint_angle, joint_velocity, joint_offset, map, pos,
                                              rot, vel, ang);
  return cudaGetLastError();
}
