// bx_kernels.cu -- instantiates the step kernel for ONE launch bound, ONE dynamics mode and with or without
// the env epilogue (compiled once per combination: -DBX_TU_MAXT=256|320|512 -DBX_TU_MINB=2|2|1 -DBX_TU_MODE=0|1
// -DBX_TU_ENV=0|1: the build runs one nvcc per object in parallel) and exports its launcher.  Variants: dynamics mode x joint class x contact class
// (bx_params.h); a system runs the variant that contains only its own features.
#include "bx_step_kernel.cuh"

#if !defined(BX_TU_MAXT) || !defined(BX_TU_MODE) || !defined(BX_TU_ENV)
#error "compile with -DBX_TU_MAXT=<threads> -DBX_TU_MINB=<blocks per SM> -DBX_TU_MODE=<0 pbd | 1 legacy_spring> -DBX_TU_ENV=<0 | 1>"
#endif

#define BX_CAT_(a, b) a##b
#define BX_CAT(a, b) BX_CAT_(a, b)
#ifndef BX_TU_NAME  // suffix of the exported launcher (default: the thread bound)
#define BX_TU_NAME BX_TU_MAXT
#endif

typedef void (*bx_kernel_fn)(const BxKernelParams);

template <int MODE, int ENV>
static bx_kernel_fn pick(int jc, int cc) {
#define BX_K(JM, CM) bx_step_kernel<MODE, (JM) | (CM) | BXF_COMMON | ENV, BX_TU_MAXT, BX_TU_MINB>
  if (MODE == BX_MODE_PBD && cc == BX_CC_PLANE_FUSED) {
    if (jc == BX_JC_REVOLUTE) return BX_K(BXF_JC0, BXF_CC0 | BXF_FUSED_CONTACTS);
    if (jc == BX_JC_SPHERICAL) return BX_K(BXF_JC1, BXF_CC0 | BXF_FUSED_CONTACTS);
    if (jc == BX_JC_SPHERICAL_ANGLE) return BX_K(BXF_JC3, BXF_CC0 | BXF_FUSED_CONTACTS);
    return BX_K(BXF_JC2, BXF_CC0 | BXF_FUSED_CONTACTS);
  }
  if (cc != BX_CC_ALL && cc != BX_CC_MESH) {
    if (jc == BX_JC_REVOLUTE) return BX_K(BXF_JC0, BXF_CC0);
    if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL) return BX_K(BXF_JC1, BXF_CC0);
    if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL_ANGLE) return BX_K(BXF_JC3, BXF_CC0);
    return BX_K(BXF_JC2, BXF_CC0);
  }
  if (cc == BX_CC_MESH) {
    // the capsule-box class is instantiated in two translation units only (320 threads x 2 tiles,
    // 512 x 1: -DBX_TU_MESH, brax_b200/build.py); bx_api.cu's launch_variant routes it there
#ifdef BX_TU_MESH
    if (jc == BX_JC_REVOLUTE) return BX_K(BXF_JC0, BXF_CC2);
    if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL) return BX_K(BXF_JC1, BXF_CC2);
    if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL_ANGLE) return BX_K(BXF_JC3, BXF_CC2);
    return BX_K(BXF_JC2, BXF_CC2);
#else
    return nullptr;
#endif
  }
  if (jc == BX_JC_REVOLUTE) return BX_K(BXF_JC0, BXF_CC1);
  if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL) return BX_K(BXF_JC1, BXF_CC1);
  if (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL_ANGLE) return BX_K(BXF_JC3, BXF_CC1);
  return BX_K(BXF_JC2, BXF_CC1);
#undef BX_K
}

static bx_kernel_fn pick_any(int mode, int jc, int cc) {
  cc &= ~BX_CC_PDL;
  if (mode != BX_TU_MODE || ((cc & BX_CC_ENV) != 0) != (BX_TU_ENV != 0)) return nullptr;   // lives in a sibling object
  if (cc & BX_CC_JO) {
    // per-env joint sockets: one all-contact-features build per (mode, joint class, env epilogue); the joint class
    // stays the system's own because the joint record's stride is a compile-time property of it (JQS)
#ifdef BX_TU_MESH
#define BX_KJ(MODE, JM, ENV) bx_step_kernel<MODE, (JM) | BXF_CC2 | BXF_COMMON | BXF_JOINT_OFFSET | (ENV), BX_TU_MAXT, BX_TU_MINB>
#define BX_KJ3(MODE, ENV)                                                                              \
  (jc == BX_JC_REVOLUTE ? BX_KJ(MODE, BXF_JC0, ENV)                                                    \
                        : (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL                                \
                               ? BX_KJ(MODE, BXF_JC1, ENV)                                             \
                               : (MODE == BX_MODE_PBD && jc == BX_JC_SPHERICAL_ANGLE ? BX_KJ(MODE, BXF_JC3, ENV) : BX_KJ(MODE, BXF_JC2, ENV))))
    return BX_KJ3(BX_TU_MODE, BX_TU_ENV ? BXF_ENV : 0);
#undef BX_KJ3
#undef BX_KJ
#else
    return nullptr;
#endif
  }
  cc &= ~BX_CC_ENV;
  return pick<BX_TU_MODE, BX_TU_ENV ? BXF_ENV : 0>(jc, cc);
}

#define BX_TU_SUFFIX BX_CAT(BX_CAT(BX_CAT(BX_CAT(BX_TU_NAME, _m), BX_TU_MODE), e), BX_TU_ENV)
cudaError_t BX_CAT(bx_launch_, BX_TU_SUFFIX)(int mode, int jc, int cc, unsigned grid, int threads,
                                           size_t smem, cudaStream_t stream,
                                           const BxKernelParams& p) {
  if (pick_any(mode, jc, cc) == nullptr) return cudaErrorInvalidDeviceFunction;
  if (!(cc & BX_CC_PDL)) {
    pick_any(mode, jc, cc)<<<grid, threads, smem, stream>>>(p);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, pick_any(mode, jc, cc), p);
}

cudaError_t BX_CAT(bx_set_smem_, BX_TU_SUFFIX)(int bytes) {
  for (int mode = BX_TU_MODE; mode <= BX_TU_MODE; mode++)
    for (int jc = 0; jc < BX_NUM_JC; jc++)
      for (int cc = 0; cc < BX_NUM_CC; cc++)
        for (int env = BX_TU_ENV; env <= BX_TU_ENV; env++)
          for (int jo = 0; jo < 2; jo++) {
            const int c2 = cc | (env ? BX_CC_ENV : 0) | (jo ? BX_CC_JO : 0);
            if (pick_any(mode, jc, c2) == nullptr) continue;
            cudaError_t e = cudaFuncSetAttribute(pick_any(mode, jc, c2), cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (e != cudaSuccess) return e;
          }
  return cudaSuccess;
}
