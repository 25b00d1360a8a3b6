// bx_ffi.cc -- XLA FFI custom-call handler over the C ABI (include/brax_b200.h).
//
// This is the JAX side of the drop-in boundary north_star names: `System.step` under jit / vmap / scan
// becomes ONE custom call `bx_step` whose handler forwards the XLA buffers to `bx_step` on XLA's stream.
// Reference call sites it serves: brax/physics/system.py:244 (System.step), brax/envs/wrappers.py:70,80
// (jp.vmap(env.step)), brax/training/acting.py:53-74 (lax.scan over env.step).
//
// Built by brax_b200/build.py::build_ffi() ONLY when jaxlib's headers resolve
// (`jax.ffi.include_dir()`); on this box jax / jaxlib are not installed and cannot be (no network), so
// the build step prints why it is skipped -- "unavailable on this box" (DESIGN.md section 7).  The
// Python registration is brax_b200/jax_ffi.py.
//
//   g++ -std=c++17 -O2 -fPIC -shared -I$(python -c "import jax.ffi; print(jax.ffi.include_dir())")
//       -I/usr/local/cuda/include -Iinclude brax_b200/csrc/bx_ffi.cc -o brax_b200/libbrax_b200_ffi.so
//       -Lbrax_b200 -lbrax_b200 -Wl,-rpath,'$ORIGIN'
#include <cuda_runtime_api.h>

#include <cstdint>

#include "../../include/brax_b200.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

using F32 = ffi::Buffer<ffi::F32>;
using R32 = ffi::ResultBuffer<ffi::F32>;

// Operands: pos[..,N,3] rot[..,N,4] vel[..,N,3] ang[..,N,3] act[..,A]; every leading (vmapped) dimension
// folds into the batch B.  Results: the 13 arrays of bx_step_io in its order.  Attribute `sys`: the
// bx_system* of brax_b200.System.handle (static tables are closed over, not traced leaves).
static ffi::Error StepImpl(cudaStream_t stream, int64_t sys_handle, F32 pos, F32 rot, F32 vel, F32 ang, F32 act,
                           R32 pos_o, R32 rot_o, R32 vel_o, R32 ang_o, R32 contact_vel, R32 contact_ang,
                           R32 joint_vel, R32 joint_ang, R32 actuator_vel, R32 actuator_ang, R32 contact_pos,
                           R32 contact_normal, R32 contact_penetration) {
  const bx_system* sys = reinterpret_cast<const bx_system*>(static_cast<intptr_t>(sys_handle));
  if (sys == nullptr) return ffi::Error::InvalidArgument("bx_step: attribute `sys` is NULL");
  const int64_t n = bx_system_num_bodies(sys), a = bx_system_act_size(sys), k = bx_system_num_contacts(sys);
  const int64_t count = static_cast<int64_t>(pos.element_count());
  if (n <= 0 || count % (3 * n) != 0) return ffi::Error::InvalidArgument("bx_step: pos is not [..., N, 3]");
  const int64_t B = count / (3 * n);
  if (static_cast<int64_t>(rot.element_count()) != 4 * n * B || static_cast<int64_t>(vel.element_count()) != 3 * n * B ||
      static_cast<int64_t>(ang.element_count()) != 3 * n * B || static_cast<int64_t>(act.element_count()) != a * B)
    return ffi::Error::InvalidArgument("bx_step: rot / vel / ang / act do not match pos");
  if (static_cast<int64_t>(contact_penetration->element_count()) != k * B)
    return ffi::Error::InvalidArgument("bx_step: contact_penetration is not [..., K]");
  bx_step_io io = {};
  io.pos = pos.typed_data(); io.rot = rot.typed_data(); io.vel = vel.typed_data(); io.ang = ang.typed_data();
  io.act = act.typed_data();
  io.pos_o = pos_o->typed_data(); io.rot_o = rot_o->typed_data(); io.vel_o = vel_o->typed_data();
  io.ang_o = ang_o->typed_data();
  io.contact_vel = contact_vel->typed_data(); io.contact_ang = contact_ang->typed_data();
  io.joint_vel = joint_vel->typed_data(); io.joint_ang = joint_ang->typed_data();
  io.actuator_vel = actuator_vel->typed_data(); io.actuator_ang = actuator_ang->typed_data();
  io.contact_pos = k ? contact_pos->typed_data() : nullptr;
  io.contact_normal = k ? contact_normal->typed_data() : nullptr;
  io.contact_penetration = k ? contact_penetration->typed_data() : nullptr;
  if (bx_step(sys, B, &io, stream) != 0) return ffi::Error::Internal(bx_last_error());
  return ffi::Error::Success();
}

XLA_FFI_DEFINE_HANDLER_SYMBOL(bx_step_ffi, StepImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Attr<int64_t>("sys")
                                  .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>()
                                  .Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>()
                                  .Ret<F32>().Ret<F32>().Ret<F32>());
