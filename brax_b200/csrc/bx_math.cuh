// bx_math.cuh -- float32 vector/quaternion helpers for the step kernel.
//
// Same formulas and the same operation order as the reference's brax/math.py
// (rotate :25-40, quat_mul :130-145, vec_quat_mul :148-170, signed_angle
// :116-127, quat_rot_axis :173-187) and brax/jumpy.py safe_norm (:170-192),
// so that the CUDA path differs from the float32 oracle only by FMA contraction.
#ifndef BX_MATH_CUH_
#define BX_MATH_CUH_

#include <cuda_runtime.h>

// BX_PACKED_V3 (default 1): component-wise V3 / Q4 arithmetic issues the packed FP32 instructions of sm_100
// (FADD2 / FMUL2 / FFMA2 = add/mul/fma.rn.f32x2: two lanes per issue slot, each rounded like the scalar
// instruction) on the (x, y) pair, the scalar one on z.  A product `v * s` / `a * b` is first an unevaluated
// proxy (V3S / V3P) so that `c + v * s` still becomes ONE fused multiply-add per component -- the contraction
// the compiler did on the scalar expression -- instead of a multiply and an add.
#ifndef BX_PACKED_V3
#define BX_PACKED_V3 1
#endif
#if BX_PACKED_V3
struct V3S;
struct V3P;
struct V3 {
  float x, y, z;
  V3() = default;
  __device__ __forceinline__ V3(const V3S& p);
  __device__ __forceinline__ V3(const V3P& p);
};
struct V3S { V3 v; float s; };   // v * s
struct V3P { V3 a, b; };         // a * b, component-wise
#else
struct V3 { float x, y, z; };
#endif
struct Q4 { float w, x, y, z; };

#define BXD __device__ __forceinline__

// Division and square root are branch-free Newton sequences on the MUFU seed
// (MUFU.RCP / MUFU.RSQ + 2 FFMA refinement + 1 residual correction).  For the
// value ranges of this path (no denormals, no overflow) the quotient / root is
// the correctly rounded one in all but a vanishing fraction of cases (<= 1 ulp
// otherwise), i.e. the same float32 result IEEE division gives, without the
// FCHK + slow-path call the compiler's a / b expands to.
// BX_FAST_DIV (default 1): a / b = a * rcp(b) with rcp = MUFU.RCP + one Newton step (<= 1 ulp):
// quotients within 1.5 ulp of the IEEE one.  0 = the correctly rounded quotient (residual
// correction, 2 more FFMA per quotient); 2 = the raw MUFU.RCP seed (A/B only).
#ifndef BX_FAST_DIV
#define BX_FAST_DIV 1
#endif
BXD float bx_rcp(float b) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
#if BX_FAST_DIV == 2
  return r;
#else
  float e = fmaf(-b, r, 1.f);
  return fmaf(r, e, r);
#endif
}
// a / b given r ~= 1/b (from bx_rcp)
BXD float bx_mul_rcp(float a, float b, float r) {
#if BX_FAST_DIV
  return a * r;
#else
  float q = a * r;
  float e = fmaf(-q, b, a);
  return fmaf(e, r, q);
#endif
}
BXD float bx_div(float a, float b) { return bx_mul_rcp(a, b, bx_rcp(b)); }
// 1 / sqrt(a), a > 0: MUFU.RSQ + one Newton step
BXD float bx_rsqrt(float a) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a));
  return y * fmaf(-0.5f * a * y, y, 1.5f);
}
BXD float bx_sqrt(float a) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a));
  float s = a * y;
  float e = fmaf(-s, s, a);
  s = fmaf(e, 0.5f * y, s);
  return a > 0.f ? s : 0.f;  // sqrt(0) = 0 (rsqrt(0) = inf); negatives never occur
}

BXD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
#if BX_PACKED_V3
BXD float2 xy(V3 a) { return make_float2(a.x, a.y); }
BXD float2 f2(float s) { return make_float2(s, s); }
BXD V3 pk(float2 r, float z) { return v3(r.x, r.y, z); }
BXD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
BXD V3 operator+(V3 a, V3 b) { return pk(__fadd2_rn(xy(a), xy(b)), a.z + b.z); }
BXD V3 operator-(V3 a, V3 b) { return pk(__fadd2_rn(xy(a), xy(-b)), a.z - b.z); }
// products: proxies, evaluated on conversion to V3
BXD V3S operator*(V3 a, float s) { V3S r; r.v = a; r.s = s; return r; }
BXD V3S operator*(float s, V3 a) { V3S r; r.v = a; r.s = s; return r; }
BXD V3P operator*(V3 a, V3 b) { V3P r; r.a = a; r.b = b; return r; }
BXD V3 ev(V3S p) { return pk(__fmul2_rn(xy(p.v), f2(p.s)), p.v.z * p.s); }
BXD V3 ev(V3P p) { return pk(__fmul2_rn(xy(p.a), xy(p.b)), p.a.z * p.b.z); }
BXD V3::V3(const V3S& p) { *this = ev(p); }
BXD V3::V3(const V3P& p) { *this = ev(p); }
BXD V3 fma3(V3 v, float s, V3 c) { return pk(__ffma2_rn(xy(v), f2(s), xy(c)), fmaf(v.z, s, c.z)); }
BXD V3 fma3(V3 a, V3 b, V3 c) { return pk(__ffma2_rn(xy(a), xy(b), xy(c)), fmaf(a.z, b.z, c.z)); }
BXD V3S operator-(V3S p) { p.s = -p.s; return p; }
BXD V3P operator-(V3P p) { p.a = -p.a; return p; }
// c + v * s and friends: one fused multiply-add per component
BXD V3 operator+(V3 c, V3S p) { return fma3(p.v, p.s, c); }
BXD V3 operator+(V3S p, V3 c) { return fma3(p.v, p.s, c); }
BXD V3 operator-(V3 c, V3S p) { return fma3(p.v, -p.s, c); }
BXD V3 operator-(V3S p, V3 c) { return fma3(p.v, p.s, -c); }
BXD V3 operator+(V3S p, V3S q) { return fma3(p.v, p.s, ev(q)); }
BXD V3 operator-(V3S p, V3S q) { return fma3(q.v, -q.s, ev(p)); }
BXD V3 operator+(V3 c, V3P p) { return fma3(p.a, p.b, c); }
BXD V3 operator+(V3P p, V3 c) { return fma3(p.a, p.b, c); }
BXD V3 operator-(V3 c, V3P p) { return fma3(-p.a, p.b, c); }
BXD V3 operator-(V3P p, V3 c) { return fma3(p.a, p.b, -c); }
BXD V3 operator+(V3P p, V3P q) { return fma3(p.a, p.b, ev(q)); }
BXD V3 operator+(V3S p, V3P q) { return fma3(p.v, p.s, ev(q)); }
BXD V3 operator+(V3P p, V3S q) { return fma3(q.v, q.s, ev(p)); }
// a product of a product: evaluate the inner one
BXD V3S operator*(V3S p, float s) { return ev(p) * s; }
BXD V3S operator*(float s, V3S p) { return ev(p) * s; }
BXD V3S operator*(V3P p, float s) { return ev(p) * s; }
BXD V3S operator*(float s, V3P p) { return ev(p) * s; }
BXD V3P operator*(V3S p, V3 b) { return ev(p) * b; }
BXD V3P operator*(V3 a, V3S p) { return a * ev(p); }
BXD V3P operator*(V3P p, V3 b) { return ev(p) * b; }
BXD V3P operator*(V3 a, V3P p) { return a * ev(p); }
#else
BXD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
BXD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
BXD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
BXD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
BXD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
BXD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
#endif
BXD V3 vdiv(V3 a, float s) {
  float r = bx_rcp(s);
  return v3(bx_mul_rcp(a.x, s, r), bx_mul_rcp(a.y, s, r), bx_mul_rcp(a.z, s, r));
}
// a / mass, given inv_mass = the float32 quotient 1 / mass from the table
#ifndef BX_INVMASS
#define BX_INVMASS 1
#endif
// Quaternion normalisation as rot * rsqrt(|rot|^2) instead of rot / sqrt(|rot|^2): OFF.  Measured: it
// breaks parity on resting bodies -- sqrt of an already unit |rot|^2 rounds to exactly 1 and the exact
// quotient returns rot unchanged, so velocity_projection (integrators.py:139-146) yields ang = 0 exactly
// for a body at rest, in the reference and here; a 1-ulp scale factor perturbs every component, leaves
// ~1e-5 rad/s of noise per substep and flips static-friction / penetration decisions.
#ifndef BX_NORM_RSQRT
#define BX_NORM_RSQRT 0
#endif
#ifndef BX_TORQUE_ARC
#define BX_TORQUE_ARC 1
#endif
BXD V3 vdivm(V3 a, float mass, float inv_mass) {
#if BX_FAST_DIV && BX_INVMASS
  return v3(a.x * inv_mass, a.y * inv_mass, a.z * inv_mass);
#else
  return v3(bx_mul_rcp(a.x, mass, inv_mass), bx_mul_rcp(a.y, mass, inv_mass), bx_mul_rcp(a.z, mass, inv_mass));
#endif
}
BXD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
BXD V3 cross(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
BXD Q4 q4(float w, float x, float y, float z) { Q4 r; r.w = w; r.x = x; r.y = y; r.z = z; return r; }
#if BX_PACKED_V3
BXD Q4 operator+(Q4 a, Q4 b) {
  const float2 l = __fadd2_rn(make_float2(a.w, a.x), make_float2(b.w, b.x)), h = __fadd2_rn(make_float2(a.y, a.z), make_float2(b.y, b.z));
  return q4(l.x, l.y, h.x, h.y);
}
BXD Q4 operator*(float s, Q4 a) {
  const float2 l = __fmul2_rn(make_float2(a.w, a.x), make_float2(s, s)), h = __fmul2_rn(make_float2(a.y, a.z), make_float2(s, s));
  return q4(l.x, l.y, h.x, h.y);
}
#else
BXD Q4 operator+(Q4 a, Q4 b) { return q4(a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z); }
BXD Q4 operator*(float s, Q4 a) { return q4(s * a.w, s * a.x, s * a.y, s * a.z); }
#endif
BXD Q4 qconj(Q4 q) { return q4(q.w, -q.x, -q.y, -q.z); }

// math.py:25-40
BXD V3 rotate(V3 v, Q4 q) {
  V3 u = v3(q.x, q.y, q.z);
  float s = q.w;
  float uv = dot(u, v), uu = dot(u, u);
  V3 r = 2.f * (uv * u) + (s * s - uu) * v;
  return r + (2.f * s) * cross(u, v);
}
// rotate((0, 0, 1), q) -- the plane normal of every plane contact row (colliders.py:672,749) -- without the terms
// that multiply the zero components: the same operations, in the same order and with the same fused
// multiply-adds `rotate` performs on the non-zero ones (adding the exact zeros changes nothing), 13 instead of ~25.
BXD V3 rotate_z(Q4 q) {
  const float uu = dot(v3(q.x, q.y, q.z), v3(q.x, q.y, q.z));
  const float d = q.w * q.w - uu, s2 = 2.f * q.w;
  const float zx = q.z * q.x, zy = q.z * q.y, zz = q.z * q.z;
  return v3(fmaf(q.y, s2, 2.f * zx), fmaf(-q.x, s2, 2.f * zy), fmaf(zz, 2.f, d));
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
// ---- inline, branch-free arctan2 / sin+cos / arccos for the substep loop ---------------
// The libm versions expand to ~70 instructions each with several reconvergence regions and
// slow-path calls (they split the basic block of a joint item, which is what the scheduler
// interleaves); these are straight-line polynomial evaluations (tools/fit_trig.py: max
// absolute error 9.5e-8 / 1.3e-7 / 2.2e-7, i.e. 1-2 ulp of the result, float32 evaluation).
#ifndef BX_LIBM_TRIG  // -DBX_LIBM_TRIG=1: the CUDA libm functions (A/B)
#define BX_LIBM_TRIG 0
#endif
// jnp.arctan2 semantics incl. (0, 0) -> 0
BXD float bx_atan2(float y, float x) {
#if BX_LIBM_TRIG
  return atan2f(y, x);
#else
  const float ax = fabsf(x), ay = fabsf(y);
  const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
  float a = mn * bx_rcp(mx);
  a = mx > 0.f ? a : 0.f;
  const float t = a * a;
  float p = 0.002903143409639597f;
  p = fmaf(p, t, -0.0162810105830431f);
  p = fmaf(p, t, 0.04303545877337456f);
  p = fmaf(p, t, -0.07533279806375504f);
  p = fmaf(p, t, 0.10654453933238983f);
  p = fmaf(p, t, -0.14207063615322113f);
  p = fmaf(p, t, 0.19993042945861816f);
  p = fmaf(p, t, -0.3333309292793274f);
  p = fmaf(p, t, 1.0f);
  float r = a * p;
  r = ay > ax ? 1.57079637050628662f - r : r;
  r = x < 0.f ? 3.14159274101257324f - r : r;
  return copysignf(r, y);
#endif
}
// sin and cos of x, |x| <= pi/2 (+2 %): half of an angle in [-pi, pi]
BXD void bx_sincos_half(float x, float* sn, float* cs) {
#if BX_LIBM_TRIG
  sincosf(x, sn, cs);
#else
  const float t = x * x;
  float ps = 2.637510533531895e-06f;
  ps = fmaf(ps, t, -0.00019824535411316901f);
  ps = fmaf(ps, t, 0.008333271369338036f);
  ps = fmaf(ps, t, -0.1666666716337204f);
  *sn = fmaf(x * t, ps, x);
  float pc = 9.446637250221102e-08f;
  pc = fmaf(pc, t, -8.698758051650657e-07f);
  pc = fmaf(pc, t, 2.61406439676648e-05f);
  pc = fmaf(pc, t, -0.0013901225756853819f);
  pc = fmaf(pc, t, 0.041667066514492035f);
  pc = fmaf(pc, t, -0.5f);
  *cs = fmaf(t, pc, 1.f);
#endif
}
// arccos(x), x in [-1, 1]
BXD float bx_acos(float x) {
#if BX_LIBM_TRIG
  return acosf(x);
#else
  const float t = fabsf(x);
  float p = -0.0012626672396436334f;
  p = fmaf(p, t, 0.006670547649264336f);
  p = fmaf(p, t, -0.01708851009607315f);
  p = fmaf(p, t, 0.030891938135027885f);
  p = fmaf(p, t, -0.05017422139644623f);
  p = fmaf(p, t, 0.08897893875837326f);
  p = fmaf(p, t, -0.21459878981113434f);
  p = fmaf(p, t, 1.570796251296997f);
  const float r = bx_sqrt(fmaxf(1.f - t, 0.f)) * p;
  return x < 0.f ? 3.14159274101257324f - r : r;
#endif
}

// ---- clip(signed_angle, lo, hi) as a DIRECTION, without arctan2 / sincos -----------------
// (x, y) = (dot(ref_p, ref_c), dot(cross(ref_p, ref_c), axis)) is r (cos psi, sin psi) of the signed
// angle psi = arctan2(y, x) (math.py:116-127).  What the joint code needs of clip(psi, lo, hi) is the
// rotation by it (joints.py:291-293, :349-352): its cosine and sine.  Inside the limits that is (x, y) / r
// -- no arctan2 -> sincos round trip, whose rounding shows up as a spurious constraint violation --
// and outside it is the limit's own (cos, sin) from the table (lcs = {cos lo, sin lo, cos hi, sin hi}).
// psi < lo / psi > hi are decided by the sign of sin(psi - limit) = (cos lim * y - sin lim * x) / r on
// the half plane (sign of y) where psi - limit cannot wrap; limits at or beyond +-pi never clip.
BXD void angle_outside(float x, float y, float4 lcs, float lo, float hi, bool* below, bool* above) {
  const float d_lo = lcs.x * y - lcs.y * x, d_hi = lcs.z * y - lcs.w * x;
  *below = lo >= 0.f ? (y < 0.f || d_lo < 0.f) : (y < 0.f && d_lo < 0.f);
  *above = hi <= 0.f ? (y > 0.f || d_hi > 0.f) : (y > 0.f && d_hi > 0.f);
}
BXD void clip_dir(float x, float y, float4 lcs, float lo, float hi, float* c, float* s, float* outside) {
  const float r2 = x * x + y * y;
  const float ri = bx_rsqrt(r2);
  bool below, above;
  angle_outside(x, y, lcs, lo, hi, &below, &above);
  float cc = r2 > 0.f ? x * ri : 1.f, ss = r2 > 0.f ? y * ri : 0.f;  // arctan2(0, 0) = 0
  cc = below ? lcs.x : cc; ss = below ? lcs.y : ss;
  cc = above ? lcs.z : cc; ss = above ? lcs.w : ss;                  // clip = min(max(psi, lo), hi)
  *c = cc; *s = ss;
  *outside = (below || above) ? 1.f : 0.f;
}
// rotate(v, quat_rot_axis(axis, phi)) for a unit axis, given cos / sin of phi (Rodrigues)
BXD V3 rot_axis_cs(V3 v, V3 axis, float c, float s) {
  return v * c + cross(axis, v) * s + axis * (dot(axis, v) * (1.f - c));
}

// math.py:116-127
BXD float signed_angle(V3 axis, V3 ref_p, V3 ref_c) {
  return bx_atan2(dot(cross(ref_p, ref_c), axis), dot(ref_p, ref_c));
}
// math.py:173-187.  The step kernel only ever passes clipped signed angles (|angle| <= pi).
BXD Q4 quat_rot_axis(V3 axis, float angle) {
  float s, c;
  bx_sincos_half(angle * 0.5f, &s, &c);
  return q4(c, axis.x * s, axis.y * s, axis.z * s);
}
// any angle (bx_reset.cu: joint angles of a reset are not range-limited)
BXD Q4 quat_rot_axis_any(V3 axis, float angle) {
  float s, c;
  sincosf(angle * 0.5f, &s, &c);
  return q4(c, axis.x * s, axis.y * s, axis.z * s);
}
// The rotation of `rotate` as a matrix (rows): 25 operations to build, 9 per vector instead of
// ~30 -- for items that rotate several constants by the same quaternion.  Same linear map as
// math.py:25-40 also for a non-unit q (it scales by |q|^2 exactly like the formula does).
#if BX_PACKED_V3
// stored by COLUMNS: m * v = c0 * v.x + c1 * v.y + c2 * v.z is three packed multiply-adds on (x, y) + three scalar
// ones on z (6 instructions instead of 9), each component the same fma(m_i2, v.z, fma(m_i1, v.y, m_i0 * v.x)) the
// row dot products contract to
struct M3 { V3 c0, c1, c2; };
#else
struct M3 { V3 r0, r1, r2; };
#endif
BXD M3 rotmat(Q4 q) {
  const float d = q.w * q.w - (q.x * q.x + q.y * q.y + q.z * q.z);
  const float x2 = q.x + q.x, y2 = q.y + q.y, z2 = q.z + q.z;
  const float xx = q.x * x2, yy = q.y * y2, zz = q.z * z2;
  const float xy = q.x * y2, xz = q.x * z2, yz = q.y * z2;
  const float sx = q.w * x2, sy = q.w * y2, sz = q.w * z2;
  M3 m;
#if BX_PACKED_V3
  m.c0 = v3(d + xx, xy + sz, xz - sy);
  m.c1 = v3(xy - sz, d + yy, yz + sx);
  m.c2 = v3(xz + sy, yz - sx, d + zz);
#else
  m.r0 = v3(d + xx, xy - sz, xz + sy);
  m.r1 = v3(xy + sz, d + yy, yz - sx);
  m.r2 = v3(xz - sy, yz + sx, d + zz);
#endif
  return m;
}
#if BX_PACKED_V3
BXD V3 mrot(const M3& m, V3 v) { return (m.c0 * v.x + m.c1 * v.y) + m.c2 * v.z; }
#else
BXD V3 mrot(const M3& m, V3 v) { return v3(dot(m.r0, v), dot(m.r1, v), dot(m.r2, v)); }
#endif
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
