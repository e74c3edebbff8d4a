// K5 -- the clip compiler on the device (see dm_clip_compiler.h).  Compile with -fmad=false.
//
// Stages, one launch each, all float64:
//   convert_frames_kernel   humanoid3d DeepMimic rows only: y-up -> z-up alignment, per-joint intrinsic x-y-z Euler
//                           angles in MJCF joint order (mocap_v2.py:56-142; mocap_util.py:31-48)
//   continuity_kernel       the shoulder / hip "continuity" fix (mocap_v2.py:143-234).  Sequential in the frame
//                           (each frame's window hangs off the previous FIXED value) but independent per joint and
//                           embarrassingly parallel over the 8 x 8 x 8 candidates: one CTA per ball joint, one
//                           thread per candidate, a (error, index) arg-min per frame that needs fixing
//   keyframe_kernel         finite-difference velocities (mocap_v2.py:274-289, calc_rot_vel :350-362) and the FK of
//                           every key frame (mocap_v2.py:292-307), one thread per key frame
//   resample_kernel         x k linear blend to the 60 Hz simulator step (mocap_v2.py:309-336), one thread per element
//   pack_rows_kernel        fp32 reward rows + state rows the step / reset kernels read (dm_step.cuh RowLayout)
#include "dm_clip_compiler.h"

#include "dm_fk_table.cuh"
#include "dm_step.cuh"
#include "generated/robot_humanoid3d.cuh"
#include "generated/robot_unitree_g1.cuh"

namespace dm {
namespace {

constexpr double kPi = 3.141592653589793;
constexpr double kEps4 = 2.220446049250313e-16 * 4.0;  // transformations.py _EPS
constexpr double kSimDt = 0.01666;                     // mocap_v2.py:311

struct Q4 {  // w, x, y, z
  double w, x, y, z;
};

__device__ __forceinline__ Q4 qmul(const Q4& a, const Q4& b) {  // Hamilton product, left-to-right sums as numpy does
  return Q4{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}

// mocap_util.py:31-40: Quaternion(matrix = Rx(+90 deg)) * q * Quaternion(matrix = Rx(-90 deg)); pyquaternion's
// trace method gives (2, +-2, 0, 0) * (0.5 / sqrt(2))
__device__ __forceinline__ Q4 align_rotation(const Q4& q) {
  const double h = 2.0 * (0.5 / sqrt(2.0));
  return qmul(qmul(Q4{h, h, 0.0, 0.0}, q), Q4{h, -h, 0.0, 0.0});
}

// euler_from_quaternion(xyzw, 'rxyz') of the vendored transformations.py (mocap_v2.py:140-142)
__device__ void euler_rxyz(const Q4& q, double e[3]) {
  double x = q.x, y = q.y, z = q.z, w = q.w;
  const double n = x * x + y * y + z * z + w * w;
  double m00 = 1.0, m01 = 0.0, m02 = 0.0, m10 = 0.0, m11 = 1.0, m12 = 0.0, m22 = 1.0;
  if (!(n < kEps4)) {
    const double s = sqrt(2.0 / n);
    x *= s; y *= s; z *= s; w *= s;
    m00 = 1.0 - y * y - z * z; m01 = x * y - z * w; m02 = x * z + y * w;
    m10 = x * y + z * w;       m11 = 1.0 - x * x - z * z; m12 = y * z - x * w;
    m22 = 1.0 - x * x - y * y;
  }
  const double cy = sqrt(m22 * m22 + m12 * m12);
  if (cy > kEps4) {
    e[0] = -atan2(m12, m22); e[1] = -atan2(-m02, cy); e[2] = -atan2(m01, m00);
  } else {
    e[0] = -0.0; e[1] = -atan2(-m02, cy); e[2] = -atan2(-m10, m11);
  }
}

// quaternion_from_euler(ex, ey, ez, 'rxyz') -> xyzw (mocap_v2.py:216)
__device__ __forceinline__ void quat_from_euler_rxyz(double ex, double ey, double ez, double q[4]) {
  const double ai = ez / 2.0, aj = -ey / 2.0, ak = ex / 2.0;
  double si, ci, sj, cj, sk, ck;
  sincos(ai, &si, &ci);
  sincos(aj, &sj, &cj);
  sincos(ak, &sk, &ck);
  const double cc = ci * ck, cs = ci * sk, sc = si * ck, ss = si * sk;
  q[0] = cj * cs - sj * sc;
  q[1] = -(cj * ss + sj * cc);
  q[2] = cj * sc - sj * cs;
  q[3] = cj * cc + sj * ss;
}

// ---- humanoid3d column maps (mocap_util.py:5-11, mocap_v2.py:62-77) ------------------------------------------
// MJCF joint order: chest, neck, right_shoulder, right_elbow, left_shoulder, left_elbow, right_hip, right_knee,
// right_ankle, left_hip, left_knee, left_ankle
constexpr int kNJoint = 12, kRawCols = 44, kNBall = 4;
__constant__ int c_src_col[kNJoint] = {8, 12, 25, 29, 39, 43, 16, 20, 21, 30, 34, 35};  // column in the raw row
__constant__ int c_dof[kNJoint] = {3, 3, 3, 1, 3, 1, 3, 1, 3, 3, 1, 3};
__constant__ int c_dst_col[kNJoint] = {7, 10, 13, 16, 17, 20, 21, 24, 25, 28, 31, 32};  // column in qpos
__constant__ int c_ball[kNJoint] = {-1, -1, 0, -1, 1, -1, 2, -1, -1, 3, -1, -1};          // continuity-fix slot
// mocap_v2.py:148-154: (x, y, z) x (min, max) of right_shoulder, left_shoulder, right_hip, left_hip
__constant__ double c_ball_lim[kNBall][3][2] = {{{-3.14, 0.50}, {-3.14, 0.70}, {-1.50, 1.50}},
                                                {{-0.50, 3.14}, {-3.14, 0.70}, {-1.50, 1.50}},
                                                {{-1.2, 1.2}, {-2.57, 1.57}, {-1.0, 1.0}},
                                                {{-1.2, 1.2}, {-2.57, 1.57}, {-1.0, 1.0}}};
__constant__ int c_ball_col[kNBall] = {13, 17, 21, 28};

__global__ void convert_frames_kernel(int n, const double* __restrict__ frames, double* __restrict__ cfg,
                                      double* __restrict__ ballq) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n) return;
  const double* r = frames + (size_t)f * kRawCols;
  double* o = cfg + (size_t)f * 35;
  // align_position: (x, y, z) -> (x, -z, y), written as the 3x3 product the reference evaluates
  const double px = r[1], py = r[2], pz = r[3];
  o[0] = 1.0 * px + 0.0 * py + 0.0 * pz;
  o[1] = 0.0 * px + 0.0 * py + -1.0 * pz;
  o[2] = 0.0 * px + 1.0 * py + 0.0 * pz;
  const Q4 root = align_rotation(Q4{r[4], r[5], r[6], r[7]});
  o[3] = root.w; o[4] = root.x; o[5] = root.y; o[6] = root.z;
  for (int j = 0; j < kNJoint; ++j) {
    const int s = c_src_col[j], d = c_dst_col[j];
    if (c_dof[j] == 1) {
      o[d] = r[s];
      continue;
    }
    const Q4 q = align_rotation(Q4{r[s], r[s + 1], r[s + 2], r[s + 3]});
    double e[3];
    euler_rxyz(q, e);
    o[d] = e[0]; o[d + 1] = e[1]; o[d + 2] = e[2];
    const int b = c_ball[j];
    if (b >= 0) {
      double* bq = ballq + ((size_t)f * kNBall + b) * 4;
      bq[0] = q.x; bq[1] = q.y; bq[2] = q.z; bq[3] = q.w;
    }
  }
}

// ---- continuity fix -----------------------------------------------------------------------------------------------
// mocap_v2.py:196-222.  window = VMX * dt.  Candidate t = ix * 64 + iy * 8 + iz walks the reference's three nested
// loops in order, so "smallest error, then smallest t" is its strict `err < best_err` update rule.
__device__ __forceinline__ double cand_value(int i, double tgt, double prev, double lo, double hi) {
  if (i == 0) return tgt;
  if (i == 1) return prev;
  if (i == 7) return hi;                    // np.linspace pins the end point
  const double step = (hi - lo) / 5.0;      // np.linspace(lo, hi, 6): arange(6) * step + lo
  return (double)(i - 2) * step + lo;
}

__global__ void __launch_bounds__(512) continuity_kernel(int n, double window, double* __restrict__ cfg,
                                                        const double* __restrict__ ballq, int* __restrict__ n_fixed) {
  const int b = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  __shared__ double s_err[16];
  __shared__ int s_idx[16];
  __shared__ int s_best;
  const int col = c_ball_col[b];
  double prev[3] = {0.0, 0.0, 0.0};
  int fixed = 0;
  for (int f = 0; f < n; ++f) {
    double* row = cfg + (size_t)f * 35 + col;
    double e[3] = {row[0], row[1], row[2]};
    if (f == 0) { prev[0] = e[0]; prev[1] = e[1]; prev[2] = e[2]; }
    double lo[3], hi[3], tgt[3];
    bool close = true;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      lo[a] = fmax(c_ball_lim[b][a][0], prev[a] - window);
      hi[a] = fmin(c_ball_lim[b][a][1], prev[a] + window);
      tgt[a] = fmin(fmax(e[a], lo[a]), hi[a]);                             // np.clip
      close = close && (fabs(e[a] - tgt[a]) <= 1e-8 + 1e-5 * fabs(tgt[a]));  // np.allclose(euler, target)
    }
    if (!close) {  // uniform across the CTA: every thread evaluated the same numbers
      const double* q = ballq + ((size_t)f * kNBall + b) * 4;
      const int ix = t >> 6, iy = (t >> 3) & 7, iz = t & 7;
      const double cx = cand_value(ix, tgt[0], prev[0], lo[0], hi[0]);
      const double cy = cand_value(iy, tgt[1], prev[1], lo[1], hi[1]);
      const double cz = cand_value(iz, tgt[2], prev[2], lo[2], hi[2]);
      double qn[4];
      quat_from_euler_rxyz(cx, cy, cz, qn);
      double dp = 0.0, dn = 0.0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const double a = qn[c] - q[c], m = -qn[c] - q[c];
        dp = dp + a * a;
        dn = dn + m * m;
      }
      const double d = fmin(sqrt(dp), sqrt(dn));
      double err = d * d;
      if (!(err == err)) err = INFINITY;
      int idx = t;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double oe = __shfl_down_sync(0xffffffffu, err, o);
        const int oi = __shfl_down_sync(0xffffffffu, idx, o);
        if (oe < err || (oe == err && oi < idx)) { err = oe; idx = oi; }
      }
      if (lane == 0) { s_err[warp] = err; s_idx[warp] = idx; }
      __syncthreads();
      if (warp == 0) {
        err = lane < 16 ? s_err[lane] : INFINITY;
        idx = lane < 16 ? s_idx[lane] : 0x7fffffff;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
          const double oe = __shfl_down_sync(0xffffffffu, err, o);
          const int oi = __shfl_down_sync(0xffffffffu, idx, o);
          if (oe < err || (oe == err && oi < idx)) { err = oe; idx = oi; }
        }
        if (lane == 0) s_best = idx;
      }
      __syncthreads();
      const int best = s_best;
      e[0] = cand_value(best >> 6, tgt[0], prev[0], lo[0], hi[0]);
      e[1] = cand_value((best >> 3) & 7, tgt[1], prev[1], lo[1], hi[1]);
      e[2] = cand_value(best & 7, tgt[2], prev[2], lo[2], hi[2]);
      if (t == 0) { row[0] = e[0]; row[1] = e[1]; row[2] = e[2]; }
      ++fixed;
      __syncthreads();  // s_best is reused by the next frame that needs fixing
    }
    prev[0] = e[0]; prev[1] = e[1]; prev[2] = e[2];
  }
  if (t == 0 && n_fixed) n_fixed[b] = fixed;
}

// ---- velocities + FK of the key frames ----------------------------------------------------------------------------
// calc_rot_vel (mocap_v2.py:350-362) with pyquaternion 0.9.9 semantics: conj(q0) * q1, renormalised unless
// |1 - |q|^2| < 1e-14; axis = v / |v| (zeros below 1e-17); angle = 2 atan2(|v|, w) wrapped into (-pi, pi].
__device__ void rot_vel(const double* q0, const double* q1, double dt, double out[3]) {
  Q4 d = qmul(Q4{q0[0], -q0[1], -q0[2], -q0[3]}, Q4{q1[0], q1[1], q1[2], q1[3]});
  const double ss = d.w * d.w + d.x * d.x + d.y * d.y + d.z * d.z;
  if (!(fabs(1.0 - ss) < 1e-14)) {
    const double n = sqrt(ss);
    if (n > 0.0) { d.w = d.w / n; d.x = d.x / n; d.y = d.y / n; d.z = d.z / n; }
  }
  const double vn = sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
  double ax = 0.0, ay = 0.0, az = 0.0;
  if (!(vn < 1e-17)) { ax = d.x / vn; ay = d.y / vn; az = d.z / vn; }
  const double theta = 2.0 * atan2(vn, d.w);
  double ang = fmod(theta + kPi, 2.0 * kPi) - kPi;  // theta >= 0: Python's % is fmod here
  if (ang == -kPi) ang = kPi;
  const double s = ang / dt;
  out[0] = s * ax; out[1] = s * ay; out[2] = s * az;
}

template <class Rb>
__global__ void keyframe_kernel(int n, double dt, const double* __restrict__ cfg, double* __restrict__ vel,
                                double* __restrict__ body_xpos, double* __restrict__ geom_xpos) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= n) return;
  const double* b = cfg + (size_t)f * Rb::NQ;
  const double* a = cfg + (size_t)(f > 0 ? f - 1 : 0) * Rb::NQ;  // frame 0 differences itself -> zeros
  double* v = vel + (size_t)f * Rb::NV;
  for (int j = 0; j < 3; ++j) v[j] = (b[j] - a[j]) / dt;
  double w[3];
  rot_vel(a + 3, b + 3, dt, w);
  v[3] = w[0]; v[4] = w[1]; v[5] = w[2];
  for (int j = 7; j < Rb::NQ; ++j) v[j - 1] = (b[j] - a[j]) / dt;
  TableIO in{b, v};
  TableSink sink{body_xpos + (size_t)f * Rb::NBODY * 3, geom_xpos + (size_t)f * Rb::NGEOM * 3};
  sink.body(0, 0.0, 0.0, 0.0);
  for (int g = 0; g < Rb::NGEOM; ++g) sink.geom(g, 0.0, 0.0, 0.0);
  FkOut<double> o;
  Rb::template fk<double>(in, o, sink);
}

// ---- x k resampling (mocap_v2.py:309-336) -------------------------------------------------------------------------
// out[ia * k + j] = (1 - j/k) * key[ia] + (j/k) * key[ia + 1]; the last key frame is dropped.  One thread per element
// of the widest table; `width` doubles per frame.
__global__ void resample_kernel(int n_out, int k, int width, const double* __restrict__ key, double* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)n_out * width) return;
  const int f = (int)(i / width), c = (int)(i - (long long)f * width);
  const int ia = f / k, j = f - ia * k;
  const double B = (double)j * 1.0 / (double)k, A = 1.0 - B;
  out[i] = A * key[(size_t)ia * width + c] + B * key[(size_t)(ia + 1) * width + c];
}

// ---- fp32 rows for the step / reset kernels (the float64 -> fp32 packing of dm_pack.h, on the device) -------------
template <class Rb>
__global__ void pack_rows_kernel(int n_rows, int first, const double* __restrict__ qpos, const double* __restrict__ qvel,
                                 const double* __restrict__ body_xpos, const double* __restrict__ geom_xpos,
                                 float* __restrict__ rows, float* __restrict__ srows) {
  using RL = RowLayout<Rb>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_rows) return;
  const int f = first + i;
  const double* q = qpos + (size_t)f * Rb::NQ;
  const double* v = qvel + (size_t)f * Rb::NV;
  float* r = rows + (size_t)i * RL::STRIDE;
  for (int c = 0; c < RL::STRIDE; ++c) r[c] = 0.0f;
  for (int s = 0; s < Rb::NSEL; ++s) {
    r[RL::Q + s] = (float)q[Rb::sel_q(s)];
    r[RL::V + s] = (float)v[Rb::sel_q(s) - 1];
  }
  double w = q[3], x = q[4], y = q[5], z = q[6], roll, pitch;  // to_rpy of the table quaternion, deepmimic_env.py:218-219
  dm_normalize4(w, x, y, z);
  dm_roll_pitch(w, x, y, z, roll, pitch);
  r[RL::ROLL] = (float)roll;
  r[RL::PITCH] = (float)pitch;
  const double* gx = geom_xpos + (size_t)f * Rb::NGEOM * 3;
  for (int e = 0; e < 4; ++e)
    for (int a = 0; a < 3; ++a) r[RL::EE + 3 * e + a] = (float)gx[3 * Rb::ee_geom(e) + a];
  const double* bx = body_xpos + (size_t)f * Rb::NBODY * 3;
  double mtot = 0.0;
  for (int b = 0; b < Rb::NBODY; ++b) mtot += Rb::body_mass(b);
  for (int a = 0; a < 3; ++a) {  // target COM, deepmimic_env.py:235-237
    double s = 0.0;
    for (int b = 0; b < Rb::NBODY; ++b) s += Rb::body_mass(b) * bx[3 * b + a];
    r[RL::COM + a] = (float)(s / mtot);
  }
  r[RL::VXY + 0] = (float)v[0];  // combined_env.py:344
  r[RL::VXY + 1] = (float)v[1];
  float* sr = srows + (size_t)i * (Rb::NQ + Rb::NV);
  for (int j = 0; j < Rb::NQ; ++j) sr[j] = (float)q[j];
  for (int j = 0; j < Rb::NV; ++j) sr[Rb::NQ + j] = (float)v[j];
}

}  // namespace

// ---- host side ------------------------------------------------------------------------------------------------------
void free_compiled_clip(CompiledClip* c) {
  if (!c) return;
  cudaFree(c->qpos); cudaFree(c->qvel); cudaFree(c->body_xpos); cudaFree(c->geom_xpos);
  *c = CompiledClip{};
}

namespace {
struct DevBuf {  // scratch that dies with the scope
  double* p = nullptr;
  ~DevBuf() { cudaFree(p); }
  bool alloc(size_t doubles) { return cudaMalloc(&p, sizeof(double) * (doubles ? doubles : 1)) == cudaSuccess; }
  double* release() { double* r = p; p = nullptr; return r; }
};

template <class Rb>
const char* compile_impl(const double* d_frames, int n, int n_cols, int direct_qpos, int getup, double dt, CompiledClip* out,
                         long long* launches, cudaStream_t st) {
  constexpr int NQ = Rb::NQ, NV = Rb::NV, NB = Rb::NBODY * 3, NG = Rb::NGEOM * 3;
  long long nl = 0;
  DevBuf cfg, vel, bx, gx, ballq;
  if (!cfg.alloc((size_t)n * NQ) || !vel.alloc((size_t)n * NV) || !bx.alloc((size_t)n * NB) || !gx.alloc((size_t)n * NG))
    return "out of device memory";
  if (direct_qpos) {  // mocap_v2.py:271-272: data_config = Frames[:, 1:]
    if (n_cols != NQ + 1) return "direct_qpos rows must have nq + 1 columns";
    if (cudaMemcpy2DAsync(cfg.p, sizeof(double) * NQ, d_frames + 1, sizeof(double) * n_cols, sizeof(double) * NQ, n,
                          cudaMemcpyDeviceToDevice, st) != cudaSuccess)
      return "cudaMemcpy2DAsync failed";
  } else {
    if (Rb::kIsG1) return "DeepMimic-format rows are defined for humanoid3d only (mocap_v2.py:56-251)";
    if (n_cols != kRawCols) return "DeepMimic rows must have 44 columns";
    if (!ballq.alloc((size_t)n * kNBall * 4)) return "out of device memory";
    convert_frames_kernel<<<(n + 63) / 64, 64, 0, st>>>(n, d_frames, cfg.p, ballq.p);
    const double window = (getup ? 5.0 : 10.0) * dt;  // VMX * dt, mocap_v2.py:156-158,197
    continuity_kernel<<<kNBall, 512, 0, st>>>(n, window, cfg.p, ballq.p, nullptr);
    nl += 2;
  }
  keyframe_kernel<Rb><<<(n + 63) / 64, 64, 0, st>>>(n, dt, cfg.p, vel.p, bx.p, gx.p);
  ++nl;
  const double ratio = dt / kSimDt;
  const int k = (int)ratio;
  if (fabs(ratio - (double)k) > 0.1) return "Invalid dt ratio, cannot interpolate mocap frames";  // mocap_v2.py:315-316
  CompiledClip c;
  c.n_key = n;
  c.k = k > 1 ? k : 1;
  if (k > 1) {
    const int len = (n - 1) * k;
    if (len <= 0) return "clip too short to resample";
    DevBuf o[4];
    double* key[4] = {cfg.p, vel.p, bx.p, gx.p};
    const int width[4] = {NQ, NV, NB, NG};
    for (int t = 0; t < 4; ++t) {
      if (!o[t].alloc((size_t)len * width[t])) return "out of device memory";
      const long long total = (long long)len * width[t];
      resample_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(len, k, width[t], key[t], o[t].p);
      ++nl;
    }
    c.len = len;
    c.dt = kSimDt;
    if (cudaStreamSynchronize(st) != cudaSuccess) return cudaGetErrorString(cudaGetLastError());
    c.qpos = o[0].release(); c.qvel = o[1].release(); c.body_xpos = o[2].release(); c.geom_xpos = o[3].release();
  } else {
    c.len = n;
    c.dt = dt;
    if (cudaStreamSynchronize(st) != cudaSuccess) return cudaGetErrorString(cudaGetLastError());
    c.qpos = cfg.release(); c.qvel = vel.release(); c.body_xpos = bx.release(); c.geom_xpos = gx.release();
  }
  if (cudaGetLastError() != cudaSuccess) { free_compiled_clip(&c); return "a clip-compiler kernel failed to launch"; }
  *out = c;
  if (launches) *launches += nl;
  return nullptr;
}

template <class Rb>
const char* pack_impl(const CompiledClip& c, int first, int n_rows, int row0, float* d_rows, float* d_state_rows,
                      long long* launches, cudaStream_t st) {
  if (first < 0 || n_rows <= 0 || first + n_rows > c.len) return "pack_rows_device: row range outside the clip";
  pack_rows_kernel<Rb><<<(n_rows + 63) / 64, 64, 0, st>>>(n_rows, first, c.qpos, c.qvel, c.body_xpos, c.geom_xpos,
                                                         d_rows + (size_t)row0 * RowLayout<Rb>::STRIDE,
                                                         d_state_rows + (size_t)row0 * (Rb::NQ + Rb::NV));
  if (launches) ++*launches;
  return cudaGetLastError() == cudaSuccess ? nullptr : "pack_rows_kernel failed to launch";
}
}  // namespace

const char* compile_clip_device(int robot, const double* d_frames, int n_frames, int n_cols, int direct_qpos, int getup,
                                double dt, CompiledClip* out, long long* launches, cudaStream_t stream) {
  if (!d_frames || n_frames <= 0 || !out) return "compile_clip_device: bad argument";
  if (robot == 0) return compile_impl<Robot_humanoid3d>(d_frames, n_frames, n_cols, direct_qpos, getup, dt, out, launches, stream);
  if (robot == 1) return compile_impl<Robot_unitree_g1>(d_frames, n_frames, n_cols, direct_qpos, getup, dt, out, launches, stream);
  return "compile_clip_device: unknown robot";
}

const char* pack_rows_device(int robot, const CompiledClip& c, int first, int n_rows, int row0, float* d_rows,
                             float* d_state_rows, long long* launches, cudaStream_t stream) {
  if (robot == 0) return pack_impl<Robot_humanoid3d>(c, first, n_rows, row0, d_rows, d_state_rows, launches, stream);
  if (robot == 1) return pack_impl<Robot_unitree_g1>(c, first, n_rows, row0, d_rows, d_state_rows, launches, stream);
  return "pack_rows_device: unknown robot";
}

}  // namespace dm
