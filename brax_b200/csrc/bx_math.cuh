// bx_math.cuh -- float32 vector/quaternion helpers for the step kernel.
//
// Same formulas and the same operation order as the reference's brax/math.py
// (rotate :25-40, quat_mul :130-145, vec_quat_mul :148-170, signed_angle
// :116-127, quat_rot_axis :173-187) and brax/jumpy.py safe_norm (:170-192),
// so that the CUDA path differs from the float32 oracle only by FMA contraction.
#ifndef BX_MATH_CUH_
#define BX_MATH_CUH_

#include <cuda_runtime.h>

struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };

#define BXD __device__ __forceinline__

// Division and square root are branch-free Newton sequences on the MUFU seed
// (MUFU.RCP / MUFU.RSQ + 2 FFMA refinement + 1 residual correction).  For the
// value ranges of this path (no denormals, no overflow) the quotient / root is
// the correctly rounded one in all but a vanishing fraction of cases (<= 1 ulp
// otherwise), i.e. the same float32 result IEEE division gives, without the
// FCHK + slow-path call the compiler's a / b expands to.
BXD float bx_rcp(float b) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
  float e = fmaf(-b, r, 1.f);
  return fmaf(r, e, r);
}
// a / b given r ~= 1/b (from bx_rcp)
BXD float bx_mul_rcp(float a, float b, float r) {
  float q = a * r;
  float e = fmaf(-q, b, a);
  return fmaf(e, r, q);
}
BXD float bx_div(float a, float b) { return bx_mul_rcp(a, b, bx_rcp(b)); }
BXD float bx_sqrt(float a) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a));
  float s = a * y;
  float e = fmaf(-s, s, a);
  s = fmaf(e, 0.5f * y, s);
  return a > 0.f ? s : 0.f;  // sqrt(0) = 0 (rsqrt(0) = inf); negatives never occur
}

BXD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
BXD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
BXD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
BXD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
BXD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
BXD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
BXD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
BXD V3 vdiv(V3 a, float s) {
  float r = bx_rcp(s);
  return v3(bx_mul_rcp(a.x, s, r), bx_mul_rcp(a.y, s, r), bx_mul_rcp(a.z, s, r));
}
BXD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
BXD V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
BXD Q4 q4(float w, float x, float y, float z) { Q4 r; r.w = w; r.x = x; r.y = y; r.z = z; return r; }
BXD Q4 operator+(Q4 a, Q4 b) { return q4(a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z); }
BXD Q4 operator*(float s, Q4 a) { return q4(s * a.w, s * a.x, s * a.y, s * a.z); }
BXD Q4 qconj(Q4 q) { return q4(q.w, -q.x, -q.y, -q.z); }

// math.py:25-40
BXD V3 rotate(V3 v, Q4 q) {
  V3 u = v3(q.x, q.y, q.z);
  float s = q.w;
  float uv = dot(u, v), uu = dot(u, u);
  V3 r = 2.f * (uv * u) + (s * s - uu) * v;
  return r + (2.f * s) * cross(u, v);
}
// math.py:130-145
BXD Q4 quat_mul(Q4 u, Q4 v) {
  return q4(u.w * v.w - u.x * v.x - u.y * v.y - u.z * v.z,
            u.w * v.x + u.x * v.w + u.y * v.z - u.z * v.y,
            u.w * v.y - u.x * v.z + u.y * v.w + u.z * v.x,
            u.w * v.z + u.x * v.y - u.y * v.x + u.z * v.w);
}
// math.py:148-170
BXD Q4 vec_quat_mul(V3 u, Q4 v) {
  return q4(-u.x * v.x - u.y * v.y - u.z * v.z,
            u.x * v.w + u.y * v.z - u.z * v.y,
            -u.x * v.z + u.y * v.w + u.z * v.x,
            u.x * v.y - u.y * v.x + u.z * v.w);
}
// math.py:116-127
BXD float signed_angle(V3 axis, V3 ref_p, V3 ref_c) {
  return atan2f(dot(cross(ref_p, ref_c), axis), dot(ref_p, ref_c));
}
// math.py:173-187
BXD Q4 quat_rot_axis(V3 axis, float angle) {
  float s, c;
  sincosf(angle * 0.5f, &s, &c);
  return q4(c, axis.x * s, axis.y * s, axis.z * s);
}
// jumpy.py:170-192: jnp path -- 0 when every |x_i| <= 1e-8 (allclose(x, 0)).
BXD float safe_norm(V3 a) {
  bool zero = fabsf(a.x) <= 1e-8f && fabsf(a.y) <= 1e-8f && fabsf(a.z) <= 1e-8f;
  float n = bx_sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
  return zero ? 0.f : n;
}
BXD float clipf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
BXD float signf(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }
BXD bool any_nz(V3 a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f; }

#endif  // BX_MATH_CUH_
