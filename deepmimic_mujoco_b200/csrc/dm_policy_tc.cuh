// K6b: the policy MLP on the 5th-generation tensor cores (tcgen05.mma, accumulators AND activations in tensor memory).
//
// Precision.  The comparison with the reference's float64 numpy is at 1e-5, which a single fp16 / bf16 / tf32 product
// (2^-11 .. 2^-8) misses by two orders of magnitude.  Every fp32 operand is therefore split into two fp16 numbers,
// x = hi + lo with hi = fp16(x), lo = fp16(x - hi) (22 significant bits; the remainder is below 2^-22 |x|, or below the
// fp16 subnormal spacing 6e-8 in absolute terms), and a product is three MMAs into the same fp32 accumulator:
//     A B  ~=  A_hi B_hi + A_hi B_lo + A_lo B_hi                      (the dropped A_lo B_lo term is 2^-22 |A B|)
// kind::f16 runs at twice the tf32 rate, so this costs what 1.5 tf32 passes would.
//
// Operands.  B (weights, N x K) comes from shared memory, K-major, no swizzle: 8 x 8 fp16 "core matrices" of 128
// contiguous bytes, [k / 8][n][k % 8] (8-row-group stride SBO = 128 B, k-chunk stride LBO = 16 N bytes), packed on the
// host per k-step of 16 as [B_hi | B_lo] and streamed by 1-D TMA bulk copies through an 8-stage ring that
// tcgen05.commit hands back.  A (activations, 128 environments x K) comes from TENSOR MEMORY: lane = environment,
// one 32-bit column = two consecutive k.  The first version of this kernel kept A in shared memory; at N = 128 an
// SS-mode MMA pulls 8 KB of operands per 64 cycles, exactly the 128 B/clk shared memory delivers, and ncu showed the
// tensor pipe 57 % busy for 12 % of useful issue (profiles/r01_policy_*).  With A in TMEM only B crosses shared memory.
// The epilogue thread that owns environment m (TMEM lane m) reads 32 accumulator columns (tcgen05.ld), applies bias +
// tanh, splits, and writes the 16 + 16 packed hi / lo columns back IN PLACE over the accumulator it has just read
// (tcgen05.st): the layer's output becomes the next layer's A operand without leaving tensor memory.
//
// TMEM columns: D0 / H1 0..255, D1 / H2 256..383, X0 384.., head accumulator 480..511 (or 0..31 when the input is
// wider than 96, which also turns the cross-tile overlap off).
// Roles (576 threads): warps 0-15 epilogue (lane quarter = warp % 4; the four warps of a quarter take every fourth
// 32-column block), warp 16 weight producer + TMEM allocation, warp 17 MMA issuer (one elected lane).  Per tile of 128 environments:
//   issuer   : L0(i) | L1(i) | L0(i+1) | head(i)        -- L0 of the next tile runs under this tile's second epilogue
//   epilogue : epi0(i) | X0(i+1) (under L1(i)) | epi1(i) | epi2(i) -> actions
#pragma once
#include <cuda_fp16.h>

#include "dm_policy.cuh"

namespace dm {

// shared memory: [weight ring: n_stages x 8 KB][raw observation tile][bias][normaliser][mbarriers]; the ring takes
// whatever the raw tile leaves (20-23 stages): deep enough that the producer streams the whole of layer 1 (128 KB)
// ahead while the epilogue warps hold the tensor core idle, instead of fetching it at L2 latency during the MMAs
constexpr int TC_STAGE_BYTES = 8192, TC_MAX_STAGES = 24;
constexpr int TC_TAIL_BYTES = 4 * (PMAX_H0 + PMAX_H1 + PHEADC) + 4 * 2 * PMAX_IN;  // bias + normaliser
enum { TB_X0 = 0, TB_D0A, TB_D0B, TB_H1, TB_D1, TB_H2, TB_D2, TB_RAW, TB_FULL, TB_EMPTY = TB_FULL + TC_MAX_STAGES,
       TB_COUNT = TB_EMPTY + TC_MAX_STAGES };
constexpr int TC_D0 = 0, TC_D1 = 256, TC_X0 = 384, TC_D2_HIGH = 480, TC_COLS = 512;
constexpr int TC_EPI_WARPS = 16, TC_EPI_THREADS = 32 * TC_EPI_WARPS, TC_THREADS = TC_EPI_THREADS + 64, TC_SUBS = TC_EPI_WARPS / 4;
constexpr int kSmemLimit = 232448;
inline int policy_tc_raw_bytes(int in_dim) { return (PM * in_dim * 4 + 1023) / 1024 * 1024; }
constexpr int TC_OUT_BYTES = PM * PHEADC * 4;  // staging tile of the actions
inline int policy_tc_stages(int in_dim) {
  const int s = (kSmemLimit - policy_tc_raw_bytes(in_dim) - TC_TAIL_BYTES - 8 * TB_COUNT - 64 - TC_OUT_BYTES) / TC_STAGE_BYTES;
  return s < TC_MAX_STAGES ? s : TC_MAX_STAGES;
}
inline size_t policy_tc_smem_bytes(int in_dim) {
  return (size_t)policy_tc_stages(in_dim) * TC_STAGE_BYTES + policy_tc_raw_bytes(in_dim) + TC_TAIL_BYTES + 8 * TB_COUNT + 16 + TC_OUT_BYTES;
}

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

// tanh of two values with THREE special-function ops instead of four: one reciprocal of the product of the two
// denominators serves both (1 / a = b / (a b)).  exp(2x) is clamped to 2^43 (tanh = 1 - 2e-13 there) so that the product
// of two exp(2x) + 1 stays below 2^88.  The epilogue is bound by the special-function unit (4 lanes per sub-partition).
// y = 2 x log2(e), already scaled (the bias is stored pre-multiplied); only the upper side needs the clamp
__device__ __forceinline__ void policy_tanh2(float y0, float y1, float& t0, float& t1) {
  y0 = fminf(y0, 43.0f);
  y1 = fminf(y1, 43.0f);
  float e0, e1, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(y0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(y1));
  const float d0 = e0 + 1.0f, d1 = e1 + 1.0f;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d0 * d1));
  r *= -2.0f;
  t0 = fmaf(r, d1, 1.0f);
  t1 = fmaf(r, d0, 1.0f);
}
// two fp32 values -> packed fp16 pair of the high parts and of the remainders
__device__ __forceinline__ void tc_split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const __half2 hh = __floats2half2_rn(x0, x1);
  const float2 back = __half22float2(hh);
  const __half2 ll = __floats2half2_rn(x0 - back.x, x1 - back.y);
  hi = *reinterpret_cast<const uint32_t*>(&hh);
  lo = *reinterpret_cast<const uint32_t*>(&ll);
}
// A-in-TMEM MMA: D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_st32(uint32_t taddr, const uint32_t (&r)[32]) {  // 32 consecutive columns of this thread's lane
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, "
      "%19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
      "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ bool tc_elect_one() {  // one lane of a converged warp
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

constexpr float kTanhScale = 2.8853900817779268f;  // 2 log2(e)

struct PolicyTcParams {
  PolicyParams p;           // sizes, buffers, normaliser
  const uint8_t* packed16;  // layer 0 (per 128-column half) | layer 1 | head; per k-step of 16: [B_hi | B_lo], each [2][N][8] fp16
  int n_stages, raw_bytes;
  long long* trace;         // optional: per-tile event times of one CTA (tools/policy_trace.py), NULL in production
  int trace_cta;
};

__global__ void __launch_bounds__(TC_THREADS, 1) policy_mlp_tc_kernel(const __grid_constant__ PolicyTcParams q) {
  extern __shared__ __align__(1024) uint8_t tsm[];
  const PolicyParams& p = q.p;
  const int NS = q.n_stages;
  float* const raw = reinterpret_cast<float*>(tsm + NS * TC_STAGE_BYTES);
  float* const s_bias = reinterpret_cast<float*>(tsm + NS * TC_STAGE_BYTES + q.raw_bytes);
  float* const s_norm = s_bias + (PMAX_H0 + PMAX_H1 + PHEADC);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(s_norm + 2 * PMAX_IN);
  uint32_t* const s_tmem = reinterpret_cast<uint32_t*>(bars + TB_COUNT);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int my_tiles = (p.n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // >= 1: grid <= n_tiles
  const int D = p.in, halves = p.h0 / 128;
  // weight stream segments: layer 0 (`halves` passes of in_pad / 16 k-steps), layer 1 (h0 / 16 k-steps), head (8)
  const int steps[3] = {p.in_pad / 16, p.h0 / 16, 8};
  const uint32_t cbytes[3] = {8192u, 8192u, (uint32_t)PHEADC * 64u};  // [hi | lo] x [2][N][8] fp16
  const uint32_t cbase[3] = {0u, (uint32_t)(halves * steps[0]) * 8192u, (uint32_t)(halves * steps[0] + steps[1]) * 8192u};
  const bool overlap = p.in_pad <= 96;  // room for the head accumulator beside X0
  const uint32_t d2col = overlap ? TC_D2_HIGH : TC_D0;
  const uint32_t tile_bytes = (uint32_t)(PM * D * 4);
  auto tile_by_tma = [&](int t) { return p.use_tma && t < p.n_tiles && (t + 1) * PM <= p.n; };
  // the order in which weight chunks are consumed: f(layer, local tile index)
  auto schedule = [&](auto&& f) {
    f(0, 0);
    for (int it = 0; it < my_tiles; ++it) {
      f(1, it);
      if (overlap && it + 1 < my_tiles) f(0, it + 1);
      f(2, it);
      if (!overlap && it + 1 < my_tiles) f(0, it + 1);
    }
  };

  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&bars[TB_FULL + s], 1); mbar_init(&bars[TB_EMPTY + s], 1); }
    mbar_init(&bars[TB_X0], TC_EPI_THREADS); mbar_init(&bars[TB_H1], TC_EPI_THREADS); mbar_init(&bars[TB_H2], TC_EPI_THREADS);
    mbar_init(&bars[TB_D0A], 1); mbar_init(&bars[TB_D0B], 1); mbar_init(&bars[TB_D1], 1); mbar_init(&bars[TB_D2], 1);
    mbar_init(&bars[TB_RAW], 1);
    mbar_fence_init();
  }
  for (int i = tid; i < p.h0 + p.h1 + PHEADC; i += TC_THREADS) s_bias[i] = p.bias[i] * (i < p.h0 + p.h1 ? kTanhScale : 1.0f);  // hidden biases pre-scaled for tanh
  for (int i = tid; i < 2 * PMAX_IN; i += TC_THREADS) s_norm[i] = p.normalize ? p.norm[i] : (i < PMAX_IN ? 0.0f : 1.0f);
  if (warp == TC_EPI_WARPS) {  // tensor-memory allocation: one warp, all 512 columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TC_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;

  if (warp == TC_EPI_WARPS) {
    // ================= weight producer =================
    if (lane == 0) {
      long long g = 0;
      int s = 0;
      uint32_t par = 1;  // a fresh mbarrier passes a wait on parity 1
      schedule([&](int layer, int) {
        const int passes = layer == 0 ? halves : 1;
        for (int c = 0; c < passes * steps[layer]; ++c, ++g) {
          mbar_wait(&bars[TB_EMPTY + s], par);
          mbar_expect_tx(&bars[TB_FULL + s], cbytes[layer]);
          tma_load_1d(tsm + s * TC_STAGE_BYTES, q.packed16 + cbase[layer] + (size_t)c * cbytes[layer], cbytes[layer], &bars[TB_FULL + s]);
          if (++s == NS) { s = 0; par ^= 1; }
        }
      });
    }
  } else if (warp == TC_EPI_WARPS + 1) {
    // ================= MMA issuer =================
    // The whole warp walks the schedule so that every operand is warp-uniform (uniform registers feed UTCHMMA directly);
    // one elected lane issues.  A first version ran this under `if (lane == 0)`: ~55 instructions and 350 cycles per
    // k-step of divergent scalar code, three times the tensor time of the step it was feeding.
    const uint32_t ring = smem_u32(tsm);
    const bool leader = tc_elect_one();
    const uint64_t bdesc128 = tc_desc(ring, 128u * 16u, 128), bdesc_head = tc_desc(ring, (uint32_t)PHEADC * 16u, 128);
    int s = 0;
    uint32_t par = 0;
    schedule([&](int layer, int it) {
      const uint32_t ph = (uint32_t)(it & 1);
      mbar_wait(&bars[layer == 0 ? TB_X0 : layer == 1 ? TB_H1 : TB_H2], ph);
      tc_fence_after();
      if (q.trace && (int)blockIdx.x == q.trace_cta && lane == 0 && it < 64) q.trace[it * 16 + 8 + layer] = clock64();  // 8, 9, 10: issue starts
      const uint32_t id = tc_idesc(layer == 2 ? PHEADC : 128);
      const uint64_t bdesc = layer == 2 ? bdesc_head : bdesc128;
      const uint64_t lo_off = layer == 2 ? (uint64_t)(PHEADC * 2) : (uint64_t)(128 * 2);  // B_lo follows B_hi: N * 32 bytes >> 4
      const int passes = layer == 0 ? halves : 1;
      for (int pass = 0; pass < passes; ++pass) {
        const uint32_t d = tmem + (layer == 0 ? TC_D0 + 128 * pass : layer == 1 ? TC_D1 : d2col);
        // A of k-step ks: layer 0 reads X0 ([hi 8 | lo 8] columns per step); layers 1 / 2 read the in-place blocks
        // ([hi 16 | lo 16] columns per 32 k) the epilogue left over D0 / D1
        const uint32_t a0 = tmem + (layer == 0 ? TC_X0 : layer == 1 ? TC_D0 : TC_D1), lo_cols = layer == 0 ? 8u : 16u;
        for (int ks = 0; ks < steps[layer]; ++ks) {
          const uint32_t a_hi = layer == 0 ? a0 + ks * 16 : a0 + (ks >> 1) * 32 + (ks & 1) * 8;
          mbar_wait(&bars[TB_FULL + s], par);
          tc_fence_after();
          const uint64_t db = bdesc + (uint64_t)(s * (TC_STAGE_BYTES >> 4));
          if (leader) {
            tc_mma_ts(d, a_hi, db, id, ks > 0);
            tc_mma_ts(d, a_hi, db + lo_off, id, 1u);
            tc_mma_ts(d, a_hi + lo_cols, db, id, 1u);
            tc_commit(&bars[TB_EMPTY + s]);
          }
          if (++s == NS) { s = 0; par ^= 1; }
        }
        if (layer == 0 && leader) tc_commit(&bars[pass == 0 ? TB_D0A : TB_D0B]);  // first half's epilogue runs under the second half's MMAs
      }
      if (layer != 0 && leader) tc_commit(&bars[layer == 1 ? TB_D1 : TB_D2]);
      if (q.trace && (int)blockIdx.x == q.trace_cta && lane == 0 && it < 64) q.trace[it * 16 + 11 + layer] = clock64();  // 11, 12, 13: issue done
    });
  } else {
    // ================= epilogue / data warps =================
    // thread owns environment row m = TMEM lane m; the two warps of a lane quarter take alternate column blocks
    const int quarter = warp & 3, sub = warp >> 2, m = quarter * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    uint32_t raw_phase = 0;
    if (tid == 0 && tile_by_tma(blockIdx.x)) {
      mbar_expect_tx(&bars[TB_RAW], tile_bytes);
      tma_load_1d(raw, p.obs + (size_t)blockIdx.x * PM * D, tile_bytes, &bars[TB_RAW]);
    }
    auto mark = [&](int it_, int ev) {
      if (q.trace && (int)blockIdx.x == q.trace_cta && tid == 0 && it_ < 64) q.trace[it_ * 16 + ev] = clock64();
    };
    int it = 0;
    // raw observations of tile t -> normalised, split, into the X0 columns; then fetch the tile after it
    auto stage_x0 = [&](int t) {
      const int e0 = t * PM, rows = min(PM, p.n - e0);
      if (tile_by_tma(t)) {
        mbar_wait(&bars[TB_RAW], raw_phase);
        raw_phase ^= 1;
      } else {
        const float* src = p.obs + (size_t)e0 * D;
        for (int i = tid; i < PM * D; i += TC_EPI_THREADS) raw[i] = i < rows * D ? ld_stream(src + i) : 0.0f;
        asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");
      }
      mark(it, 14);
      if (p.reward_norm && tid < rows)
        p.reward_norm[e0 + tid] = dm_max_nan(dm_min_nan(p.reward[e0 + tid] * __ldg(p.norm + 2 * PMAX_IN), p.clip_r), -p.clip_r);
      float* const on = p.obs_norm && m < rows ? p.obs_norm + (size_t)(e0 + m) * D : nullptr;
      const int l0_steps = p.in_pad / 16;
      const float bound = p.normalize ? fminf(p.clip_obs, 65504.0f) : 65504.0f;
      for (int ks = sub; ks < l0_steps; ks += TC_SUBS) {
        // branch-free: 48 independent shared-memory loads first (columns past D read the neighbouring row and are
        // zeroed below; the normaliser arrays are padded with mean 0 / scale 1), arithmetic after
        const float* rp = raw + m * D + ks * 16;
        const float* np_ = s_norm + ks * 16;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = (rp[i] - np_[i]) * np_[PMAX_IN + i];
        // one clamp serves the normaliser's clip and the fp16 range
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = dm_max_nan(dm_min_nan(v[i], bound), -bound);
        if (ks == l0_steps - 1) {  // the only k-step with columns past D
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = ks * 16 + i < D ? v[i] : 0.0f;
        }
        if (on) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (ks * 16 + i < D) on[ks * 16 + i] = v[i];
        }
        uint32_t r[16];
#pragma unroll
        for (int i = 0; i < 8; ++i)
          tc_split2(v[2 * i], v[2 * i + 1], r[i], r[8 + i]);
        tc_st16(lane_base + TC_X0 + ks * 16, r);
      }
      mark(it, 15);
      tc_wait_st();
      tc_fence_before();
      mbar_arrive(&bars[TB_X0]);
      asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");  // every thread has finished reading `raw`
      const int tn = t + gridDim.x;
      if (tid == 0 && tile_by_tma(tn)) {
        fence_async_smem();
        mbar_expect_tx(&bars[TB_RAW], tile_bytes);
        tma_load_1d(raw, p.obs + (size_t)tn * PM * D, tile_bytes, &bars[TB_RAW]);
      }
    };
    // accumulator columns [c0, c0 + cols) -> tanh(. + bias) -> packed hi / lo written back over the same columns
    auto epilogue_tanh = [&](uint32_t c0, int cols, const float* bias) {
      for (int cb = 32 * sub; cb < cols; cb += 32 * TC_SUBS) {
        float v[32];
        tc_ld32(lane_base + c0 + cb, v);
        uint32_t r[32];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float t0, t1;
          policy_tanh2(fmaf(v[2 * i], kTanhScale, bias[cb + 2 * i]), fmaf(v[2 * i + 1], kTanhScale, bias[cb + 2 * i + 1]), t0, t1);
          tc_split2(t0, t1, r[i], r[16 + i]);
        }
        tc_st32(lane_base + c0 + cb, r);
      }
      tc_wait_st();
      tc_fence_before();
    };
    // head accumulator of tile t -> + bias, clip -> staging tile -> coalesced stores
    float* const out_stage = reinterpret_cast<float*>(tsm + NS * TC_STAGE_BYTES + q.raw_bytes + TC_TAIL_BYTES + 8 * TB_COUNT + 16);
    auto head_out = [&](int t, uint32_t ph, int it_) {
      const int e0 = t * PM, rows = min(PM, p.n - e0);
      mbar_wait(&bars[TB_D2], ph);
      tc_fence_after();
      mark(it_, 6);  // head ready
      constexpr int HC = PHEADC / TC_SUBS;  // head columns per warp
      float v[HC];
      tc_ld8(lane_base + d2col + HC * sub, v);
      tc_fence_before();
      const float* hb = s_bias + p.h0 + p.h1 + HC * sub;
#pragma unroll
      for (int j = 0; j < HC; ++j)
        if (HC * sub + j < p.out) out_stage[m * p.out + HC * sub + j] = dm_max_nan(dm_min_nan(v[j] + hb[j], p.act_hi), p.act_lo);
      asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");
      float* dst = p.action + (size_t)e0 * p.out;
      for (int i = tid; i < rows * p.out; i += TC_EPI_THREADS) dst[i] = out_stage[i];
      // the staging tile is rewritten one tile later, behind the barriers of stage_x0
    };
    mark(0, 7);
    stage_x0(blockIdx.x);
    int pending = -1;
    for (int t = blockIdx.x; t < p.n_tiles; t += gridDim.x, ++it) {
      const uint32_t ph = (uint32_t)(it & 1);
      const int e0 = t * PM, rows = min(PM, p.n - e0), tn = t + gridDim.x;
      for (int h = 0; h < halves; ++h) {
        mbar_wait(&bars[h == 0 ? TB_D0A : TB_D0B], ph);
        tc_fence_after();
        mark(it, h);  // 0, 1: layer-0 half ready
        epilogue_tanh(TC_D0 + 128 * h, 128, s_bias + 128 * h);
        if (h == 0 && pending >= 0) {
          head_out(pending, ph ^ 1u, it - 1);
          pending = -1;
        }
      }
      mbar_arrive(&bars[TB_H1]);
      mark(it, 2);  // H1 handed over
      if (overlap && tn < p.n_tiles) stage_x0(tn);  // X0 is free (layer 0 has retired); runs under layer 1
      mark(it, 3);  // next X0 staged
      mbar_wait(&bars[TB_D1], ph);
      tc_fence_after();
      mark(it, 4);  // layer 1 ready
      epilogue_tanh(TC_D1, 128, s_bias + p.h0);
      mbar_arrive(&bars[TB_H2]);
      mark(it, 5);  // H2 handed over
      if (!overlap) {  // the head accumulator shares columns with D0: drain it before the next layer 0 may start
        head_out(t, ph, it);
        if (tn < p.n_tiles) stage_x0(tn);
      } else {
        pending = t;   // drained after the first half of the next tile's epilogue, so the head MMAs run under it
      }
    }
    if (pending >= 0) head_out(pending, (uint32_t)((it - 1) & 1), it - 1);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TC_EPI_WARPS) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_COLS) : "memory");
}

}  // namespace dm
