// K6: observation / reward normaliser + policy MLP in one kernel, so that the observation the step kernel wrote is
// consumed on the device.  The network is the reference's ExtractedPolicy.act
// (src/extracted_policy.py:471-478; arch [256, 128], src/sb3_ppo.py:265):
//     a = tanh(tanh(x W0 + B0) W2 + B2) WA + BA,        x = clip((obs - mean) * inv_std, +-clip_obs)
// fp32 on the CUDA cores (the comparison with the reference's float64 numpy is at 1e-5; bf16 / tf32 tensor-core
// products are 1e-3).  Roofline: fp32 FMA issue -- 53.5 k FMA per environment for 66 -> 256 -> 128 -> 28 against
// 0.4 KB of HBM traffic.
//
// Launch shape: persistent CTAs (one per SM, 256 threads), tiles of 128 environments.
//   * activations live in shared memory, k-major ([feature][environment]), so one 128-bit load feeds four rows;
//   * the weights are packed ONCE on the host into 16 x 128 (head: 16 x 32) chunks in the exact order the kernel
//     consumes them, and streamed from L2 through a 3-stage ring with 1-D TMA bulk copies (mbarrier completion);
//     the stream runs on across layer, pass and tile boundaries, so the pipeline never drains;
//   * every thread owns an 8 x 8 (head: 8 x 2) block of the 128 x 128 output tile: per k step four 128-bit
//     shared-memory loads feed 64 FMAs;
//   * the next tile's raw observations are fetched by TMA during the head layer; outputs leave through a staging
//     tile with coalesced stores.
#pragma once
#include "dm_tile_kernel.cuh"

namespace dm {

constexpr int PM = 128;                  // environments per tile
constexpr int PK = 16;                   // k rows per weight chunk
constexpr int PSTAGES = 3;               // ring stages
constexpr int PCHUNK = PK * 128;         // floats per ring stage (8 KB)
constexpr int PHEADC = 32;               // padded head width
constexpr int PMAX_IN = 128, PMAX_H0 = 256, PMAX_H1 = 128, PMAX_OUT = PHEADC;
constexpr int P_H1 = 0, P_H2 = PMAX_H0 * PM, P_RING = P_H2 + PMAX_H1 * PM, P_BIAS = P_RING + PSTAGES * PCHUNK,
              P_NORM = P_BIAS + PMAX_H0 + PMAX_H1 + PHEADC, P_FLOATS = P_NORM + 2 * PMAX_IN;
constexpr int P_OUTSTAGE = PMAX_IN * PM;  // inside the H1 region, behind the raw observation tile
constexpr size_t kPolicySmemBytes = sizeof(float) * P_FLOATS + 8 * (PSTAGES + 1);

struct PolicyParams {
  int n, n_tiles;
  int in, in_pad, h0, h1, out;  // in_pad: multiple of PK
  int big_chunks, chunks_per_tile;
  const float* obs;
  float* action;
  float* obs_norm;           // optional: the normalised observations, [n, in]
  const float* reward;       // optional
  float* reward_norm;
  const float* packed;
  const float* bias;         // [h0 | h1 | PHEADC]
  const float* norm;         // [mean (PMAX_IN) | inv_std (PMAX_IN) | reward scale]
  float clip_obs, clip_r, act_lo, act_hi;
  int normalize, use_tma;
};

// tanh with ~1e-7 ABSOLUTE error: 1 - 2 / (exp(2x) + 1) on the special-function unit (the hardware tanh.approx is 5e-4)
__device__ __forceinline__ float policy_tanh(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
  return fmaf(-2.0f, r, 1.0f);
}

struct PolicyPipe {
  float* ring;
  uint64_t* full;
  const float* packed;
  int big_chunks, chunks_per_tile;
  long long total;  // chunks this CTA will consume
  long long g;      // chunks consumed so far
  int stage, phase;
  __device__ __forceinline__ void issue(long long gc) const {  // one thread
    const int c = (int)(gc % chunks_per_tile), s = (int)(gc % PSTAGES);
    const bool big = c < big_chunks;
    const float* src = packed + (big ? (size_t)c * PCHUNK : (size_t)big_chunks * PCHUNK + (size_t)(c - big_chunks) * (PK * PHEADC));
    const uint32_t bytes = big ? PCHUNK * 4u : PK * PHEADC * 4u;
    mbar_expect_tx(&full[s], bytes);
    tma_load_1d(ring + s * PCHUNK, src, bytes, &full[s]);
  }
  __device__ __forceinline__ const float* acquire() {
    mbar_wait(&full[stage], (uint32_t)phase);
    return ring + stage * PCHUNK;
  }
  __device__ __forceinline__ void release(int tid) {  // all threads; ends with the chunk's CTA barrier
    __syncthreads();
    if (tid == 0 && g + PSTAGES < total) issue(g + PSTAGES);
    ++g;
    if (++stage == PSTAGES) { stage = 0; phase ^= 1; }
  }
};

// one 128-column pass: acc[i][j] = sum_k src[k][row_i] * W[k][col_j]; rows {4 tx .. +3} U {64 + 4 tx .. +3}, cols 8 ty .. +7
__device__ __forceinline__ void policy_pass8(PolicyPipe& pp, const float* __restrict__ src, int n_chunks, int tid, float (&acc)[8][8]) {
  const int tx = tid & 15, ty = tid >> 4;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
  for (int ch = 0; ch < n_chunks; ++ch) {
    const float* Bs = pp.acquire() + ty * 8;
    const float* As = src + (size_t)ch * PK * PM + tx * 4;
#pragma unroll
    for (int kk = 0; kk < PK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(As + kk * PM);
      const float4 a1 = *reinterpret_cast<const float4*>(As + kk * PM + 64);
      const float4 b0 = *reinterpret_cast<const float4*>(Bs + kk * 128);
      const float4 b1 = *reinterpret_cast<const float4*>(Bs + kk * 128 + 4);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    pp.release(tid);
  }
}

// dst[col][row] = tanh(acc + bias[col]) for the pass's 128 columns starting at n0
__device__ __forceinline__ void policy_store_tanh(float* __restrict__ dst, const float* __restrict__ bias, int n0, int tid,
                                                  const float (&acc)[8][8]) {
  const int tx = tid & 15, ty = tid >> 4;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int col = n0 + ty * 8 + j;
    const float bj = bias[col];
    float* d = dst + (size_t)col * PM + tx * 4;
    *reinterpret_cast<float4*>(d) = make_float4(policy_tanh(acc[0][j] + bj), policy_tanh(acc[1][j] + bj),
                                                policy_tanh(acc[2][j] + bj), policy_tanh(acc[3][j] + bj));
    *reinterpret_cast<float4*>(d + 64) = make_float4(policy_tanh(acc[4][j] + bj), policy_tanh(acc[5][j] + bj),
                                                     policy_tanh(acc[6][j] + bj), policy_tanh(acc[7][j] + bj));
  }
}

__global__ void __launch_bounds__(256, 1) policy_mlp_kernel(const __grid_constant__ PolicyParams p) {
  extern __shared__ __align__(128) float psm[];
  float* const H1 = psm + P_H1;
  float* const H2 = psm + P_H2;   // also X0, the normalised input tile, k-major
  float* const s_bias = psm + P_BIAS;
  float* const s_norm = psm + P_NORM;
  uint64_t* const bars = reinterpret_cast<uint64_t*>(psm + P_FLOATS);  // [PSTAGES] weight stages, [PSTAGES] raw obs
  const int tid = threadIdx.x;
  const int my_tiles = (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (my_tiles <= 0) return;
  PolicyPipe pp{psm + P_RING, bars, p.packed, p.big_chunks, p.chunks_per_tile, (long long)my_tiles * p.chunks_per_tile, 0, 0, 0};
  uint64_t* const raw_bar = bars + PSTAGES;
  const int D = p.in;
  const uint32_t tile_bytes = (uint32_t)(PM * D * 4);
  if (tid == 0) {
    for (int s = 0; s <= PSTAGES; ++s) mbar_init(&bars[s], 1);
    mbar_fence_init();
  }
  for (int i = tid; i < p.h0 + p.h1 + PHEADC; i += 256) s_bias[i] = p.bias[i];
  for (int i = tid; i < 2 * PMAX_IN; i += 256) s_norm[i] = p.normalize ? p.norm[i] : (i < PMAX_IN ? 0.0f : 1.0f);
  __syncthreads();
  auto tile_by_tma = [&](int t) { return p.use_tma && (t + 1) * PM <= p.n; };
  if (tid == 0) {
    for (int c = 0; c < PSTAGES && c < pp.total; ++c) pp.issue(c);
    if (tile_by_tma(blockIdx.x)) {
      mbar_expect_tx(raw_bar, tile_bytes);
      tma_load_1d(H1, p.obs + (size_t)blockIdx.x * PM * D, tile_bytes, raw_bar);
    }
  }
  uint32_t raw_phase = 0;
  const int l0_chunks = p.in_pad / PK, l1_chunks = p.h0 / PK, l2_chunks = p.h1 / PK;
  for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x) {
    const int e0 = t * PM, rows = min(PM, p.n - e0);
    // ---- raw observation tile -> H1[0 .. 128 D) ----
    if (tile_by_tma(t)) {
      mbar_wait(raw_bar, raw_phase);
      raw_phase ^= 1;
    } else {
      const float* src = p.obs + (size_t)e0 * D;
      for (int i = tid; i < PM * D; i += 256) H1[i] = i < rows * D ? ld_stream(src + i) : 0.0f;
      __syncthreads();
    }
    // ---- normalise in place (row-major, coalesced write of obs_norm), then transpose into X0 ----
    if (p.normalize || p.obs_norm) {
      for (int i = tid; i < rows * D; i += 256) {
        const int k = i % D;
        float x = (H1[i] - s_norm[k]) * s_norm[PMAX_IN + k];
        if (p.normalize) x = dm_max_nan(dm_min_nan(x, p.clip_obs), -p.clip_obs);
        H1[i] = x;
        if (p.obs_norm) p.obs_norm[(size_t)e0 * D + i] = x;
      }
      __syncthreads();
    }
    if (p.reward_norm && tid < rows)
      p.reward_norm[e0 + tid] = dm_max_nan(dm_min_nan(p.reward[e0 + tid] * __ldg(p.norm + 2 * PMAX_IN), p.clip_r), -p.clip_r);
    {
      const int m = tid & (PM - 1);
      for (int k = tid >> 7; k < p.in_pad; k += 2) H2[(size_t)k * PM + m] = k < D ? H1[m * D + k] : 0.0f;
    }
    __syncthreads();
    float acc[8][8];
    // ---- layer 0: X0 (H2 region) -> H1 ----
    for (int n0 = 0; n0 < p.h0; n0 += 128) {
      policy_pass8(pp, H2, l0_chunks, tid, acc);
      policy_store_tanh(H1, s_bias, n0, tid, acc);
    }
    __syncthreads();
    // ---- layer 1: H1 -> H2 ----
    policy_pass8(pp, H1, l1_chunks, tid, acc);
    policy_store_tanh(H2, s_bias + p.h0, 0, tid, acc);
    __syncthreads();
    // H1 is dead: fetch the next tile's raw observations behind the head layer
    const int tn = t + gridDim.x;
    if (tid == 0 && tn < p.n_tiles && tile_by_tma(tn)) {
      fence_async_smem();
      mbar_expect_tx(raw_bar, tile_bytes);
      tma_load_1d(H1, p.obs + (size_t)tn * PM * D, tile_bytes, raw_bar);
    }
    // ---- head: H2 -> out staging; 8 rows x 2 columns per thread ----
    {
      const int tx = tid & 15, ty = tid >> 4;
      float h[8][2];
#pragma unroll
      for (int i = 0; i < 8; ++i) h[i][0] = h[i][1] = 0.0f;
      for (int ch = 0; ch < l2_chunks; ++ch) {
        const float* Bs = pp.acquire() + ty * 2;
        const float* As = H2 + (size_t)ch * PK * PM + tx * 4;
#pragma unroll
        for (int kk = 0; kk < PK; ++kk) {
          const float4 a0 = *reinterpret_cast<const float4*>(As + kk * PM);
          const float4 a1 = *reinterpret_cast<const float4*>(As + kk * PM + 64);
          const float2 b = *reinterpret_cast<const float2*>(Bs + kk * PHEADC);
          const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            h[i][0] = fmaf(a[i], b.x, h[i][0]);
            h[i][1] = fmaf(a[i], b.y, h[i][1]);
          }
        }
        pp.release(tid);
      }
      float* so = H1 + P_OUTSTAGE;
      const float* hb = s_bias + p.h0 + p.h1;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = ty * 2 + j;
        if (col < p.out) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int row = (i < 4 ? 0 : 60) + tx * 4 + i;
            so[row * p.out + col] = dm_max_nan(dm_min_nan(h[i][j] + hb[col], p.act_hi), p.act_lo);
          }
        }
      }
    }
    __syncthreads();
    {
      const float* so = H1 + P_OUTSTAGE;
      float* dst = p.action + (size_t)e0 * p.out;
      for (int i = tid; i < rows * p.out; i += 256) dst[i] = so[i];
    }
    // the staging tile is rewritten only after the many barriers of the next tile's layers
  }
}

// mean / inv_std of the observation columns and the reward scale from the step kernels' running sums (dm_stats)
__global__ void policy_norm_from_stats_kernel(const double* __restrict__ stats, int D, float eps, float* __restrict__ norm,
                                              float* __restrict__ r_scale) {  // r_scale = norm + 2 PMAX_IN
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const double n = stats[0] > 1.0 ? stats[0] : 1.0;
  if (k < D) {
    const double mean = stats[6 + k] / n;
    double var = stats[6 + D + k] / n - mean * mean;
    var = var > 0.0 ? var : 0.0;
    norm[k] = (float)mean;
    norm[PMAX_IN + k] = (float)(1.0 / sqrt(var + (double)eps));
  }
  if (k == 0) {
    const double rm = stats[1] / n;
    double rv = stats[2] / n - rm * rm;
    rv = rv > 0.0 ? rv : 0.0;
    *r_scale = (float)(1.0 / sqrt(rv + (double)eps));
  }
}

}  // namespace dm
