// Scalar helpers shared by the generated forward kinematics and the step logic.
// Everything is templated on the scalar type: float in the per-step kernel, double in the
// clip-table builder.  The file also compiles as plain C++ (tests/host_harness builds it with
// g++ to unit-test the arithmetic on machines without a GPU); the shipped library is CUDA only.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define DM_HD __host__ __device__ __forceinline__
#else
#define DM_HD inline
#endif

namespace dm {

template <typename T>
struct FkOut {
  T rq[4];        // root (free joint) quaternion after mj_normalizeQuat, wxyz
  T tq[4];        // torso body xquat
  T com[3];       // sum(m * xipos) / sum(m)   (mass_center / subtree_com of the root)
  T com_xpos[3];  // sum(m * xpos) / sum(m)    (the reward's "C.O.M", deepmimic_env.py:235-238)
  T ee[12];       // 4 end-effector geom_xpos, RobotConfig.endeffector_geom_names order
  T cvel[6];      // torso cvel: [angular; linear]
  uint32_t contact;  // floor contacts derived from FK (bit layout of the dm_step_args contact word)
};

struct NullSink {
  static constexpr bool kWantsGeoms = false;
  DM_HD static constexpr bool contact_on() { return false; }  // floor-contact tests of the FK (DM_CONTACT_FROM_FK)
  template <typename T> DM_HD void body(int, T, T, T) {}
  template <typename T> DM_HD void geom(int, T, T, T) {}
};

DM_HD float dm_abs(float a) { return fabsf(a); }
DM_HD double dm_abs(double a) { return fabs(a); }
DM_HD float dm_fma(float a, float b, float c) { return fmaf(a, b, c); }
DM_HD double dm_fma(double a, double b, double c) { return fma(a, b, c); }

DM_HD uint32_t dm_f2u(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
DM_HD float dm_u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// sin and cos, branch-free and short (the step spends a quarter of its instructions here: one call per hinge):
// quadrant k = round(x * 2/pi) by the magic-number add (k lands in the low mantissa bits, no F2I / FRND),
// two-term FMA reduction by pi/2 (each step is a single rounding of an exact product-sum, so the reduced
// argument is good to ~3e-8 for |x| <= 200), the Cephes single-precision minimax polynomials on
// [-pi/4, pi/4], quadrant swap by two selects and the signs by integer XOR.  Abs. error < 1.5e-7; any joint
// angle beyond 100 rad trips the observation-bounds guard, which discards the step anyway.
DM_HD void dm_sincos_f(float x, float& s, float& c) {
  const float t = fmaf(x, 0.63661977236758134f, 12582912.0f);  // 1.5 * 2^23
  const uint32_t k = dm_f2u(t);                                // low bits = round(x * 2/pi) (two's complement)
  const float kf = t - 12582912.0f;
  float r = fmaf(kf, -1.5707963705062866f, x);
  r = fmaf(kf, 4.371138828673793e-8f, r);
  const float r2 = r * r;
  float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
  ps = fmaf(ps, r2, -1.6666654611e-1f);
  ps = fmaf(ps * r2, r, r);
  float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
  pc = fmaf(pc, r2, 4.166664568298827e-2f);
  pc = fmaf(pc * r2, r2, fmaf(r2, -0.5f, 1.0f));
  const bool swap = (k & 1u) != 0;
  const float ss = swap ? pc : ps, cc = swap ? ps : pc;
  s = dm_u2f(dm_f2u(ss) ^ ((k << 30) & 0x80000000u));
  c = dm_u2f(dm_f2u(cc) ^ (((k + 1u) << 30) & 0x80000000u));
}
DM_HD void dm_sincos_half(float theta, float& s, float& c) { dm_sincos_f(0.5f * theta, s, c); }

// Two hinges at once.  On the device the fp32 version runs the argument reduction and both polynomials of BOTH angles in
// Blackwell's packed fp32 arithmetic (fma / mul / add .rn.f32x2, SASS FFMA2 / FMUL2 / FADD2: one issue slot for two lanes,
// constants as broadcast immediates), lane for lane the same IEEE operations as dm_sincos_f, so the results are bit-identical
// to two scalar calls: 14 packed + 4 constant moves + 14 integer / select instructions for the pair against 2 x 25.  The
// generated FK pairs consecutive hinges (tools/gen_model_code.py).
#if defined(__CUDACC__)
__device__ __forceinline__ unsigned long long dm_pk2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void dm_upk2(unsigned long long v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long dm_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ unsigned long long dm_mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long dm_add2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ void dm_sincos_sign(uint32_t k, float ps, float pc, float& s, float& c) {
  const bool swap = (k & 1u) != 0;
  const float ss = swap ? pc : ps, cc = swap ? ps : pc;
  s = dm_u2f(dm_f2u(ss) ^ ((k << 30) & 0x80000000u));
  c = dm_u2f(dm_f2u(cc) ^ (((k + 1u) << 30) & 0x80000000u));
}
#endif
DM_HD void dm_sincos_half2(float ta, float tb, float& sa, float& ca, float& sb, float& cb) {
#if defined(__CUDA_ARCH__)
#define DM_K2(v) dm_pk2(v, v)
  const unsigned long long x = dm_mul2(dm_pk2(ta, tb), DM_K2(0.5f));
  const unsigned long long t = dm_fma2(x, DM_K2(0.63661977236758134f), DM_K2(12582912.0f));
  const unsigned long long kf = dm_add2(t, DM_K2(-12582912.0f));
  unsigned long long r = dm_fma2(kf, DM_K2(-1.5707963705062866f), x);
  r = dm_fma2(kf, DM_K2(4.371138828673793e-8f), r);
  const unsigned long long r2 = dm_mul2(r, r);
  unsigned long long ps = dm_fma2(r2, DM_K2(-1.9515295891e-4f), DM_K2(8.3321608736e-3f));
  ps = dm_fma2(ps, r2, DM_K2(-1.6666654611e-1f));
  ps = dm_fma2(dm_mul2(ps, r2), r, r);
  unsigned long long pc = dm_fma2(r2, DM_K2(2.443315711809948e-5f), DM_K2(-1.388731625493765e-3f));
  pc = dm_fma2(pc, r2, DM_K2(4.166664568298827e-2f));
  pc = dm_fma2(dm_mul2(pc, r2), r2, dm_fma2(r2, DM_K2(-0.5f), DM_K2(1.0f)));
#undef DM_K2
  float t0, t1, ps0, ps1, pc0, pc1;
  dm_upk2(t, t0, t1);
  dm_upk2(ps, ps0, ps1);
  dm_upk2(pc, pc0, pc1);
  dm_sincos_sign(dm_f2u(t0), ps0, pc0, sa, ca);
  dm_sincos_sign(dm_f2u(t1), ps1, pc1, sb, cb);
#else
  dm_sincos_half(ta, sa, ca);
  dm_sincos_half(tb, sb, cb);
#endif
}
DM_HD void dm_sincos_half(double theta, double& s, double& c) {
  s = sin(0.5 * theta);
  c = cos(0.5 * theta);
}
DM_HD void dm_sincos_half2(double ta, double tb, double& sa, double& ca, double& sb, double& cb) {
  dm_sincos_half(ta, sa, ca);
  dm_sincos_half(tb, sb, cb);
}

DM_HD float dm_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
  return rsqrtf(x);
#else
  return 1.0f / sqrtf(x);
#endif
}
DM_HD double dm_rsqrt(double x) { return 1.0 / sqrt(x); }

DM_HD float dm_rcp(float x) {  // the ONE consumer is dm_atan2_f's ratio in [0, 1]: MUFU.RCP (1 ulp) adds <= 1.2e-7 to the angle,
                               // the correctly rounded __frcp_rn is 13 instructions with a slow-path branch
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}

// atan2f, branch-free and short: a = min(|x|,|y|) / max(|x|,|y|) in [0, 1], odd minimax polynomial
// a + a s q(s), s = a^2 (fit error 8e-9, fp32 evaluation ~1e-7 abs), octant fix-ups by selects, sign by XOR.
// atan2(+-0, +-0) = +-0 / +-pi like libm; NaN in -> NaN out.  CUDA's atan2f is ~40 instructions with a slow-path
// branch; the step calls it five times.
DM_HD float dm_atan2_f(float y, float x) {
  const float ax = fabsf(x), ay = fabsf(y);
  const float mx = ax > ay ? ax : ay, mn = ax > ay ? ay : ax;
  float a = mn * dm_rcp(mx);
  a = mx == 0.0f ? 0.0f : a;                       // atan2(0, 0)
  a = (ax == ay && ax > 3.0e38f) ? 1.0f : a;       // both infinite
  const float s2 = a * a;
  float q = fmaf(s2, 2.658708952e-03f, -1.528460532e-02f);
  q = fmaf(q, s2, 4.138148949e-02f);
  q = fmaf(q, s2, -7.390017062e-02f);
  q = fmaf(q, s2, 1.058567539e-01f);
  q = fmaf(q, s2, -1.418923885e-01f);
  q = fmaf(q, s2, 1.999084800e-01f);
  q = fmaf(q, s2, -3.333300948e-01f);
  float r = fmaf(a * s2, q, a);
  r = ay > ax ? 1.57079637f - r : r;
  r = (dm_f2u(x) & 0x80000000u) ? 3.14159274f - r : r;
  return dm_u2f(dm_f2u(r) | (dm_f2u(y) & 0x80000000u));  // r >= 0; a NaN input has already made r NaN
}

// mju_normalize4: zero quaternion -> identity
template <typename T>
DM_HD void dm_normalize4(T& w, T& x, T& y, T& z) {
  T n2 = w * w + x * x + y * y + z * z;
  if (!(n2 >= T(1e-30))) {  // mjMINVAL^2; also catches NaN
    if (n2 == n2) { w = T(1); x = T(0); y = T(0); z = T(0); }
    return;
  }
  T inv = T(1) / sqrt(n2);
  w *= inv; x *= inv; y *= inv; z *= inv;
}

// mju_quat2Mat
template <typename T>
DM_HD void dm_quat2mat(T w, T x, T y, T z, T& r00, T& r01, T& r02, T& r10, T& r11, T& r12, T& r20, T& r21, T& r22) {
  T ww = w * w, xx = x * x, yy = y * y, zz = z * z;
  T xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
  r00 = ww + xx - yy - zz; r01 = T(2) * (xy - wz);   r02 = T(2) * (xz + wy);
  r10 = T(2) * (xy + wz);  r11 = ww - xx + yy - zz;  r12 = T(2) * (yz - wx);
  r20 = T(2) * (xz - wy);  r21 = T(2) * (yz + wx);   r22 = ww - xx - yy + zz;
}
// single columns of the same matrix (world direction of a body-frame coordinate axis)
template <typename T>
DM_HD void dm_quat_col0(T w, T x, T y, T z, T& a, T& b, T& c) {
  a = w * w + x * x - y * y - z * z; b = T(2) * (x * y + w * z); c = T(2) * (x * z - w * y);
}
template <typename T>
DM_HD void dm_quat_col1(T w, T x, T y, T z, T& a, T& b, T& c) {
  a = T(2) * (x * y - w * z); b = w * w - x * x + y * y - z * z; c = T(2) * (y * z + w * x);
}
template <typename T>
DM_HD void dm_quat_col2(T w, T x, T y, T z, T& a, T& b, T& c) {
  a = T(2) * (x * z + w * y); b = T(2) * (y * z - w * x); c = w * w - x * x - y * y + z * z;
}

// q <- q * (c, s*e_k): hinge about a coordinate axis (mju_axisAngle2Quat + mju_mulQuat)
template <typename T>
DM_HD void dm_quat_hinge_x(T& w, T& x, T& y, T& z, T s, T c) {
  T nw = w * c - x * s, nx = x * c + w * s, ny = y * c + z * s, nz = z * c - y * s;
  w = nw; x = nx; y = ny; z = nz;
}
template <typename T>
DM_HD void dm_quat_hinge_y(T& w, T& x, T& y, T& z, T s, T c) {
  T nw = w * c - y * s, nx = x * c - z * s, ny = y * c + w * s, nz = z * c + x * s;
  w = nw; x = nx; y = ny; z = nz;
}
template <typename T>
DM_HD void dm_quat_hinge_z(T& w, T& x, T& y, T& z, T s, T c) {
  T nw = w * c - z * s, nx = x * c + y * s, ny = y * c - x * s, nz = z * c + w * s;
  w = nw; x = nx; y = ny; z = nz;
}

#if defined(__CUDA_ARCH__)
// fp32 device overloads in packed arithmetic: the quaternion travels as the pairs P = (w, x), Q = (y, z); ptxas folds the
// half swaps, the per-half signs and the scalar broadcasts below into FFMA2 / FMUL2 operand modifiers (.LO_HI, .NP, .F32),
// so a hinge is 2 FMUL2 + 2 FFMA2 instead of 4 FMUL + 4 FFMA.  (mju_quat2Mat in the same pairing was tried twice: nothing for
// humanoid3d -- the scalar form only computes the entries a body's children use, and its products already contract into
// FMAs -- and for unitree_g1's 38 full matrices 104 instructions fewer but SLOWER: G1 combined SoA, 1 M envs 250 -> 267 us.)
__device__ __forceinline__ void dm_quat_hinge_x(float& w, float& x, float& y, float& z, float s, float c) {
  const unsigned long long C = dm_pk2(c, c);
  const unsigned long long nP = dm_fma2(dm_pk2(x, w), dm_pk2(-s, s), dm_mul2(dm_pk2(w, x), C));
  const unsigned long long nQ = dm_fma2(dm_pk2(z, y), dm_pk2(s, -s), dm_mul2(dm_pk2(y, z), C));
  dm_upk2(nP, w, x);
  dm_upk2(nQ, y, z);
}
__device__ __forceinline__ void dm_quat_hinge_y(float& w, float& x, float& y, float& z, float s, float c) {
  const unsigned long long C = dm_pk2(c, c), P = dm_pk2(w, x), Q = dm_pk2(y, z);
  const unsigned long long nP = dm_fma2(Q, dm_pk2(-s, -s), dm_mul2(P, C));
  const unsigned long long nQ = dm_fma2(P, dm_pk2(s, s), dm_mul2(Q, C));
  dm_upk2(nP, w, x);
  dm_upk2(nQ, y, z);
}
__device__ __forceinline__ void dm_quat_hinge_z(float& w, float& x, float& y, float& z, float s, float c) {
  const unsigned long long C = dm_pk2(c, c);
  const unsigned long long nP = dm_fma2(dm_pk2(z, y), dm_pk2(-s, s), dm_mul2(dm_pk2(w, x), C));
  const unsigned long long nQ = dm_fma2(dm_pk2(x, w), dm_pk2(-s, s), dm_mul2(dm_pk2(y, z), C));
  dm_upk2(nP, w, x);
  dm_upk2(nQ, y, z);
}
#endif

// py3dtf Quaternion(x, y, z, w).to_rpy() of a UNIT quaternion (see oracle/rotations.py:to_rpy)
template <typename T>
DM_HD void dm_rpy(T w, T x, T y, T z, T& roll, T& pitch, T& yaw) {
  roll = atan2(T(2) * (w * x + y * z), T(1) - T(2) * (x * x + y * y));
  T sp = T(2) * (w * y - z * x);
  sp = sp > T(1) ? T(1) : (sp < T(-1) ? T(-1) : sp);
  pitch = asin(sp);
  yaw = atan2(T(2) * (w * z + x * y), T(1) - T(2) * (y * y + z * z));
}
template <typename T>
DM_HD void dm_roll_pitch(T w, T x, T y, T z, T& roll, T& pitch) {
  roll = atan2(T(2) * (w * x + y * z), T(1) - T(2) * (x * x + y * y));
  T sp = T(2) * (w * y - z * x);
  sp = sp > T(1) ? T(1) : (sp < T(-1) ? T(-1) : sp);
  pitch = asin(sp);
}

// float overload of dm_rpy (per-step path): the short atan2, asin(x) = atan2(x, sqrt(1 - x^2))
DM_HD void dm_rpy(float w, float x, float y, float z, float& roll, float& pitch, float& yaw) {
  roll = dm_atan2_f(2.0f * (w * x + y * z), 1.0f - 2.0f * (x * x + y * y));
  float sp = 2.0f * (w * y - z * x);
  sp = sp > 1.0f ? 1.0f : (sp < -1.0f ? -1.0f : sp);
  pitch = dm_atan2_f(sp, sqrtf(fmaf(-sp, sp, 1.0f)));
  yaw = dm_atan2_f(2.0f * (w * z + x * y), 1.0f - 2.0f * (y * y + z * z));
}

// Roll / pitch / yaw of the RAW (fp32, not necessarily unit) free-joint quaternion.  Products of fp32
// values are exact in double, so the atan2 arguments carry no cancellation error; that keeps roll and
// yaw accurate next to gimbal lock (the face-down get-up poses sit at pitch ~ 90 deg), where the fp32
// forms 1 - 2(x^2 + y^2) lose all their digits.  pitch uses atan2(sp, cp), which stays well conditioned.
DM_HD void dm_rpy_raw(float wf, float xf, float yf, float zf, float& roll, float& pitch, float& yaw) {
  const double w = wf, x = xf, y = yf, z = zf;
  const double ww = w * w, xx = x * x, yy = y * y, zz = z * z;
  const double n2 = (ww + xx) + (yy + zz);
  const double rn = 2.0 * (w * x + y * z), rd = n2 - 2.0 * (xx + yy);
  const double yn = 2.0 * (w * z + x * y), yd = n2 - 2.0 * (yy + zz);
  const double sp = 2.0 * (w * y - z * x);
  const double cp2 = yn * yn + yd * yd;  // = (n2 cos(pitch))^2
  roll = dm_atan2_f((float)rn, (float)rd);
  yaw = dm_atan2_f((float)yn, (float)yd);
  pitch = dm_atan2_f((float)sp, sqrtf((float)cp2));
  if (!(n2 >= 1e-30)) {  // mju_normalize4 turns a zero quaternion into identity
    if (n2 == n2) { roll = 0.0f; pitch = 0.0f; yaw = 0.0f; }
  }
}

// numpy max / min semantics: NaN propagates (np.max(obs) > 100 is False when any obs is NaN)
DM_HD float dm_max_nan(float a, float b) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
#else
  return (a != a || b != b) ? NAN : (a > b ? a : b);
#endif
}
DM_HD float dm_min_nan(float a, float b) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
#else
  return (a != a || b != b) ? NAN : (a < b ? a : b);
#endif
}

}  // namespace dm
