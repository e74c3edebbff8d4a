// K6b: the policy MLP on the 5th-generation tensor cores (tcgen05.mma, accumulators in tensor memory).
//
// Precision.  The comparison with the reference's float64 numpy is at 1e-5, which a single fp16 / bf16 / tf32 product
// (2^-11 .. 2^-8) misses by two orders of magnitude.  Every fp32 operand is therefore split into two fp16 numbers,
// x = hi + lo with hi = fp16(x), lo = fp16(x - hi) (22 significant bits; the remainder is below 2^-22 |x|, or below the
// fp16 subnormal spacing 6e-8 in absolute terms), and a product is three MMAs into the same fp32 accumulator:
//     A B  ~=  A_hi B_hi + A_hi B_lo + A_lo B_hi                      (the dropped A_lo B_lo term is 2^-22 |A B|)
// kind::f16 runs at twice the tf32 rate, so this costs what 1.5 tf32 passes would and needs no more shared memory than
// fp32 operands (2 + 2 bytes per element).
//
// Operands.  A (activations, 128 environments x K) and B (weights, N x K) are K-major, no swizzle: 8 x 8 fp16 "core
// matrices" of 128 contiguous bytes, [k / 8][row][k % 8], so the 8-row-group stride (SBO) is 128 B and the k-chunk
// stride (LBO) is rows * 16 B.  An epilogue thread owns one environment (one TMEM lane): eight consecutive outputs of a
// layer are exactly one 16-byte row of a core matrix of the next layer's A operand -- one conflict-free 128-bit store
// per eight values, no transposition.
//
// Roles (192 threads): warps 0-3 = 128 epilogue threads (TMEM lane quarter = warp id), warp 4 = weight producer (1-D TMA
// bulk copies through an 8-stage ring, released by tcgen05.commit) + TMEM allocation, warp 5 = MMA issuer (one lane).
// Per tile of 128 environments, with layer 0 / layer 1 split into two 128-wide halves so that activations fit beside
// the ring (R0 = X0 then H2, R1 = raw observations then one half of H1; 64 KB each):
//   issuer   : L0a L0b | L1a (k < 128) | L1b (k >= 128) | head
//   epilogue : X0 | epi0a (D0a -> H1 half) | epi0b (after L1a has consumed the first half) | epi1 (D1 -> H2) | epi2 -> out
// The next tile's raw observations are fetched by TMA as soon as L1b has retired.
#pragma once
#include <cuda_fp16.h>

#include "dm_policy.cuh"

namespace dm {

constexpr int TC_STAGES = 8, TC_STAGE_BYTES = 8192;
constexpr int TC_R0 = 0, TC_R1 = 65536, TC_RING = 131072, TC_BIAS = TC_RING + TC_STAGES * TC_STAGE_BYTES,
              TC_NORM = TC_BIAS + 4 * (PMAX_H0 + PMAX_H1 + PHEADC), TC_BARS = TC_NORM + 4 * 2 * PMAX_IN;
enum { TB_FULL = 0, TB_EMPTY = TC_STAGES, TB_X0 = 2 * TC_STAGES, TB_D0A, TB_D0B, TB_H1A, TB_H1FREE, TB_H1B, TB_D1, TB_H2, TB_D2,
       TB_RAW, TB_COUNT };
constexpr size_t kPolicyTcSmemBytes = TC_BARS + 8 * TB_COUNT + 16;
constexpr int TC_HALF = 32768;  // hi | lo split of a 64 KB region
constexpr int TC_D0 = 0, TC_D1 = 256, TC_D2 = 384, TC_COLS = 512;

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {  // arrives on `bar` when all MMAs issued so far have retired
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// shared-memory matrix descriptor: K-major, no swizzle (layout type 0), descriptor version 1 (Blackwell)
__device__ __forceinline__ uint64_t tc_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor: D fp32, A / B fp16, both K-major, M = 128
__device__ __forceinline__ constexpr uint32_t tc_idesc(int n) { return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24); }
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint64_t a, uint64_t b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a), "l"(b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float (&v)[32]) {  // 32 consecutive columns of this thread's lane
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
      "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// eight fp32 values -> one core-matrix row of the hi array and one of the lo array
__device__ __forceinline__ void tc_split_store(uint8_t* hi_row, uint8_t* lo_row, const float (&x)[8]) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half2 hh = __floats2half2_rn(x[2 * i], x[2 * i + 1]);
    const float2 back = __half22float2(hh);
    const __half2 ll = __floats2half2_rn(x[2 * i] - back.x, x[2 * i + 1] - back.y);
    h[i] = *reinterpret_cast<const uint32_t*>(&hh);
    l[i] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  *reinterpret_cast<uint4*>(hi_row) = make_uint4(h[0], h[1], h[2], h[3]);
  *reinterpret_cast<uint4*>(lo_row) = make_uint4(l[0], l[1], l[2], l[3]);
}

struct PolicyTcParams {
  PolicyParams p;          // sizes, buffers, normaliser (packed / big_chunks / chunks_per_tile describe the fp16 stream)
  const uint8_t* packed16; // per k-step of 16: [B_hi | B_lo], each [2][N][8] fp16 (N = 128, head: 32)
};

__global__ void __launch_bounds__(192, 1) policy_mlp_tc_kernel(const __grid_constant__ PolicyTcParams q) {
  extern __shared__ __align__(1024) uint8_t tsm[];
  const PolicyParams& p = q.p;
  uint8_t* const R0 = tsm + TC_R0;
  uint8_t* const R1 = tsm + TC_R1;
  float* const s_bias = reinterpret_cast<float*>(tsm + TC_BIAS);
  float* const s_norm = reinterpret_cast<float*>(tsm + TC_NORM);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(tsm + TC_BARS);
  uint32_t* const s_tmem = reinterpret_cast<uint32_t*>(bars + TB_COUNT);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int my_tiles = (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // >= 1: grid <= n_tiles
  const int D = p.in, halves = p.h0 / 128, l0_steps = p.in_pad / 16;
  const uint32_t tile_bytes = (uint32_t)(PM * D * 4);
  auto tile_by_tma = [&](int t) { return p.use_tma && (t + 1) * PM <= p.n; };

  if (tid == 0) {
    for (int s = 0; s < TC_STAGES; ++s) { mbar_init(&bars[TB_FULL + s], 1); mbar_init(&bars[TB_EMPTY + s], 1); }
    mbar_init(&bars[TB_X0], 128); mbar_init(&bars[TB_H1A], 128); mbar_init(&bars[TB_H1B], 128); mbar_init(&bars[TB_H2], 128);
    mbar_init(&bars[TB_D0A], 1); mbar_init(&bars[TB_D0B], 1); mbar_init(&bars[TB_H1FREE], 1); mbar_init(&bars[TB_D1], 1);
    mbar_init(&bars[TB_D2], 1); mbar_init(&bars[TB_RAW], 1);
    mbar_fence_init();
  }
  for (int i = tid; i < p.h0 + p.h1 + PHEADC; i += 192) s_bias[i] = p.bias[i];
  for (int i = tid; i < 2 * PMAX_IN; i += 192) s_norm[i] = p.normalize ? p.norm[i] : (i < PMAX_IN ? 0.0f : 1.0f);
  if (warp == 4) {  // tensor-memory allocation: one warp, all 512 columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TC_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;

  if (warp == 4) {
    // ================= weight producer =================
    if (lane == 0) {
      const long long total = (long long)my_tiles * p.chunks_per_tile;
      for (long long g = 0; g < total; ++g) {
        const int s = (int)(g % TC_STAGES), c = (int)(g % p.chunks_per_tile);
        mbar_wait(&bars[TB_EMPTY + s], (uint32_t)(((g / TC_STAGES) & 1) ^ 1));
        const bool big = c < p.big_chunks;
        const uint8_t* src = q.packed16 + (big ? (size_t)c * 8192 : (size_t)p.big_chunks * 8192 + (size_t)(c - p.big_chunks) * 2048);
        const uint32_t bytes = big ? 8192u : 2048u;
        mbar_expect_tx(&bars[TB_FULL + s], bytes);
        tma_load_1d(tsm + TC_RING + s * TC_STAGE_BYTES, src, bytes, &bars[TB_FULL + s]);
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t r0 = smem_u32(R0), r1 = smem_u32(R1), ring = smem_u32(tsm + TC_RING);
      long long g = 0;
      // one k-step of 16: D (+)= A_hi B_hi + A_hi B_lo + A_lo B_hi, then hand the stage back
      auto kstep = [&](uint32_t d, uint32_t a_hi, uint32_t a_lo, int n, uint32_t acc) {
        const int s = (int)(g % TC_STAGES);
        mbar_wait(&bars[TB_FULL + s], (uint32_t)((g / TC_STAGES) & 1));
        tc_fence_after();
        const uint32_t b = ring + s * TC_STAGE_BYTES, half = (uint32_t)n * 32u;
        const uint64_t da_hi = tc_desc(a_hi, 2048, 128), da_lo = tc_desc(a_lo, 2048, 128);
        const uint64_t db_hi = tc_desc(b, (uint32_t)n * 16u, 128), db_lo = tc_desc(b + half, (uint32_t)n * 16u, 128);
        const uint32_t id = tc_idesc(n);
        tc_mma(d, da_hi, db_hi, id, acc);
        tc_mma(d, da_hi, db_lo, id, 1u);
        tc_mma(d, da_lo, db_hi, id, 1u);
        tc_commit(&bars[TB_EMPTY + s]);
        ++g;
      };
      for (int it = 0; it < my_tiles; ++it) {
        const uint32_t ph = (uint32_t)(it & 1);
        mbar_wait(&bars[TB_X0], ph);
        tc_fence_after();
        for (int h = 0; h < halves; ++h) {  // layer 0: X0 (R0) -> D0 half h
          for (int ks = 0; ks < l0_steps; ++ks) kstep(tmem + TC_D0 + 128 * h, r0 + ks * 4096, r0 + TC_HALF + ks * 4096, 128, ks > 0);
          tc_commit(&bars[h == 0 ? TB_D0A : TB_D0B]);
        }
        for (int h = 0; h < halves; ++h) {  // layer 1: H1 half h (R1) -> D1
          mbar_wait(&bars[h == 0 ? TB_H1A : TB_H1B], ph);
          tc_fence_after();
          for (int ks = 0; ks < 8; ++ks) kstep(tmem + TC_D1, r1 + ks * 4096, r1 + TC_HALF + ks * 4096, 128, (h | ks) > 0);
          tc_commit(&bars[h == halves - 1 ? TB_D1 : TB_H1FREE]);
        }
        mbar_wait(&bars[TB_H2], ph);  // head: H2 (R0) -> D2
        tc_fence_after();
        for (int ks = 0; ks < 8; ++ks) kstep(tmem + TC_D2, r0 + ks * 4096, r0 + TC_HALF + ks * 4096, PHEADC, ks > 0);
        tc_commit(&bars[TB_D2]);
      }
    }
  } else {
    // ================= epilogue / data warps: thread m owns environment row m and TMEM lane m =================
    const int m = tid;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    float* const raw = reinterpret_cast<float*>(R1);
    uint32_t raw_phase = 0;
    if (tid == 0 && tile_by_tma(blockIdx.x)) {
      mbar_expect_tx(&bars[TB_RAW], tile_bytes);
      tma_load_1d(raw, p.obs + (size_t)blockIdx.x * PM * D, tile_bytes, &bars[TB_RAW]);
    }
    // D -> tanh(D + bias) -> next layer's A operand (hi at dst, lo at dst + TC_HALF), `cols` columns starting at TMEM column c0
    auto epilogue_tanh = [&](uint32_t c0, int cols, const float* bias, uint8_t* dst) {
      for (int cb = 0; cb < cols; cb += 32) {
        float v[32];
        tc_ld32(lane_base + c0 + cb, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float x[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) x[i] = policy_tanh(v[8 * j + i] + bias[cb + 8 * j + i]);
          uint8_t* row = dst + ((cb >> 3) + j) * 2048 + m * 16;
          tc_split_store(row, row + TC_HALF, x);
        }
      }
      fence_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
      tc_fence_before();
    };
    int it = 0;
    for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x, ++it) {
      const uint32_t ph = (uint32_t)(it & 1);
      const int e0 = t * PM, rows = min(PM, p.n - e0);
      // ---- raw observation tile in R1 ----
      if (tile_by_tma(t)) {
        mbar_wait(&bars[TB_RAW], raw_phase);
        raw_phase ^= 1;
      } else {
        const float* src = p.obs + (size_t)e0 * D;
        for (int i = tid; i < PM * D; i += 128) raw[i] = i < rows * D ? ld_stream(src + i) : 0.0f;
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      if (p.normalize || p.obs_norm) {
        for (int i = tid; i < rows * D; i += 128) {
          const int k = i % D;
          float x = (raw[i] - s_norm[k]) * s_norm[PMAX_IN + k];
          if (p.normalize) x = dm_max_nan(dm_min_nan(x, p.clip_obs), -p.clip_obs);
          raw[i] = x;
          if (p.obs_norm) p.obs_norm[(size_t)e0 * D + i] = x;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      if (p.reward_norm && tid < rows)
        p.reward_norm[e0 + tid] = dm_max_nan(dm_min_nan(p.reward[e0 + tid] * __ldg(p.norm + 2 * PMAX_IN), p.clip_r), -p.clip_r);
      // ---- X0: this thread's row, split, into R0 ----
      for (int kc = 0; kc < p.in_pad / 8; ++kc) {
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int k = kc * 8 + i;
          const float r = k < D ? raw[m * D + k] : 0.0f;
          x[i] = fminf(fmaxf(r, -65504.0f), 65504.0f);  // fp16 range (NaN -> -65504: the guard upstream zeroes such rows)
        }
        uint8_t* row = R0 + kc * 2048 + m * 16;
        tc_split_store(row, row + TC_HALF, x);
      }
      fence_async_smem();
      mbar_arrive(&bars[TB_X0]);
      // ---- layer 0 halves -> H1 half (R1) ----
      for (int h = 0; h < halves; ++h) {
        mbar_wait(&bars[h == 0 ? TB_D0A : TB_D0B], ph);
        if (h == 1) mbar_wait(&bars[TB_H1FREE], ph);  // layer 1 has consumed the first half
        tc_fence_after();
        epilogue_tanh(TC_D0 + 128 * h, 128, s_bias + 128 * h, R1);
        mbar_arrive(&bars[h == 0 ? TB_H1A : TB_H1B]);
      }
      // ---- layer 1 -> H2 (R0) ----
      mbar_wait(&bars[TB_D1], ph);
      tc_fence_after();
      {  // R1 is dead: fetch the next tile's raw observations now
        const int tn = t + gridDim.x;
        if (tid == 0 && tn < p.n_tiles && tile_by_tma(tn)) {
          fence_async_smem();
          mbar_expect_tx(&bars[TB_RAW], tile_bytes);
          tma_load_1d(raw, p.obs + (size_t)tn * PM * D, tile_bytes, &bars[TB_RAW]);
        }
      }
      epilogue_tanh(TC_D1, 128, s_bias + p.h0, R0);
      mbar_arrive(&bars[TB_H2]);
      // ---- head -> actions ----
      mbar_wait(&bars[TB_D2], ph);
      tc_fence_after();
      {
        float v[32];
        tc_ld32(lane_base + TC_D2, v);
        tc_fence_before();
        if (m < rows) {
          const float* hb = s_bias + p.h0 + p.h1;
          float* dst = p.action + (size_t)(e0 + m) * p.out;
#pragma unroll
          for (int j = 0; j < PHEADC; ++j)
            if (j < p.out) dst[j] = dm_max_nan(dm_min_nan(v[j] + hb[j], p.act_hi), p.act_lo);
        }
      }
      if (!tile_by_tma(t + gridDim.x)) asm volatile("bar.sync 1, 128;" ::: "memory");  // cooperative raw load of the next tile overwrites R1
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_COLS) : "memory");
}

}  // namespace dm
