// K1b: row-major (AoS) step kernel -- persistent CTAs, shared-memory tiles, TMA in, in-place tile, TMA out.
//
// Launch shape: one CTA per SM, G groups of 128 threads per CTA.  Each group walks tiles of TE = 128 consecutive
// environments (tile = cta * G + group, stride grid * G) and synchronises only with itself (named barriers), so
// one group's TMA wait overlaps the others' arithmetic.  The four warps of a group sit on the four SM
// sub-partitions and run in step; groups run out of phase.  Both halves of that matter and were measured
// (profiles/r01_design_notes.md): the per-environment step is ~2500 straight-line instructions (40-100 KB of
// SASS, beyond the 32 KB L1.5 instruction cache), so 16 fully autonomous warps each stream the whole kernel from
// L2 (1.3 "no instruction" stall cycles per issue, 48 % issue utilisation), while a single CTA-wide lock-step
// tile stalls every warp at once on each global-load latency (42 %).  Four groups of four keep the instruction
// streams at four and still hide latency.
// Per tile:
//   in    the tile's qpos and qvel rows are contiguous in the row-major arrays: two 1-D TMA bulk loads
//         (cp.async.bulk global -> shared, completion on the group's mbarrier; SASS: UBLKCP); the group's next
//         tile is pulled into L2 as soon as this one has landed (cp.async.bulk.prefetch.L2), the first one already in
//         front of griddepcontrol.wait;
//   table the packed clip table is staged ONCE per CTA in shared memory (when it fits), so the per-environment
//         gather of a mocap frame is a 128-bit shared-memory load;
//   step  every thread steps its own environment out of shared memory; the computed observation entries stay
//         in registers;
//   out   the observation tile [TE, obs_dim] is assembled IN PLACE: every thread first pulls its joint blocks
//         into registers, the group syncs, every thread writes its finished row over the (now dead) input rows
//         -- the row stride obs_dim is odd, or 2 mod 32 at worst, so the stores are (nearly) conflict-free --
//         and one thread issues a single TMA bulk store shared -> global for the whole tile.  Nothing is
//         written to global memory with a row stride, and an observation entry costs ~2 shared-memory
//         instructions instead of a trip through a cooperative gather loop;
//   torque action -> torque is a flat 128-bit pass global -> global issued while the TMA load is in flight;
//   stats the normaliser moments are column sums of the finished tile read back from shared memory (lanes walk
//         columns, warps walk rows: conflict-free); five per-environment scalars (1, reward, done, episode
//         return and length of finished episodes) ride along as extra columns.  The per-thread fp64 running sums
//         live in shared memory, not registers, and are flushed with one CTA-level reduction and one set of fp64
//         atomics per CTA.
// Batches that are not a multiple of 128 environments hand the remainder to the strided kernel (K1a).
#pragma once
#include "dm_kernels.cuh"
#include "dm_peer.cuh"

namespace dm {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared (16-byte aligned, size multiple of 16), completes on `bar`
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// 1-D TMA bulk copy shared -> global, tracked by the issuing thread's bulk async-group
__device__ __forceinline__ void tma_store_1d(void* dst, const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all bulk stores of this thread have finished READING shared memory (the buffer may be overwritten)
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// generic-proxy writes to shared memory become visible to the async proxy (TMA)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void l2_prefetch(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

// streaming global loads that bypass L1 allocation: the tile kernel leaves the SM only ~30-60 KB of L1, which the
// register spills (local memory) need more than one-touch inputs do
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
template <class T>
__device__ __forceinline__ T ld_stream(const T* p) {
  static_assert(sizeof(T) == 4, "32-bit");
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
  return *reinterpret_cast<T*>(&v);
}

template <int KIND, class Rb>
struct TileCfg {
  using OL = ObsLayout<KIND, Rb>;
#ifndef DM_TE
#define DM_TE 128
#endif
  static constexpr int TE = DM_TE;                     // environments per tile = threads per group
  static constexpr int NW = TE / 32;                   // warps per group
  static constexpr int NCOMP = OL::DIM - 2 * Rb::NJ;  // computed obs entries (torso, contacts, phase, player action)
  // a tile buffer row: the input rows (nq + nv floats), later the observation row plus one statistics column (reward)
  static constexpr int ROW = (Rb::NQ + Rb::NV) > OL::DIM + 1 ? (Rb::NQ + Rb::NV) : OL::DIM + 1;
  static constexpr int NCOLS = OL::DIM + 1, NCOLIT = (NCOLS + 31) / 32;
  // robots whose action rows are not a whole number of 128-bit words / actuators (unitree_g1: 23 actions, 37 motors) get
  // their action tile staged behind the rows by a third bulk copy; the torque pass then reads shared memory instead of
  // stalling twice per tile on global loads (ncu: 30 % of the G1 kernel's stall samples sat on that pass)
  static constexpr bool kVecTorque = Rb::NA == Rb::NU && Rb::NA % 4 == 0;
  static constexpr int ACT_FLOATS = kVecTorque ? 0 : TE * Rb::NA;
  static_assert((ACT_FLOATS * 4) % 16 == 0, "bulk copies move multiples of 16 bytes");
  static constexpr int BUF_FLOATS = TE * ROW + ACT_FLOATS;
  static constexpr size_t BUF_BYTES = sizeof(float) * BUF_FLOATS;              // one ring slot
  static constexpr size_t ACC_BYTES = sizeof(double) * (2 * NCOLIT * TE + 4);  // per group, STATS only
  static constexpr int MAXBUF = 12;                                            // ring slots (mbarriers 0 .. MAXBUF-1, table = MAXBUF)
  static constexpr int NUP = (Rb::NU + 3) / 4 * 4;
  static constexpr size_t FIXED_BYTES = 128 + sizeof(float) * 3 * NUP + 64;  // mbarriers + motor table (gear | lo | hi) + issued[16]
  static_assert(FIXED_BYTES % 16 == 0 && BUF_BYTES % 128 == 0, "alignment");
  static constexpr size_t SMEM_BUDGET = 227 * 1024;
};

// per-thread accessor: inputs from the staged rows, computed observation entries into registers
template <int KIND, class Rb>
struct TileIO {
  static constexpr bool kZeroOnGuard = false;  // the kernel zeroes the row when it writes the tile
  static constexpr int NCOMP = TileCfg<KIND, Rb>::NCOMP;
  const float* __restrict__ qp;  // this env's qpos row in shared memory
  const float* __restrict__ qv;
  float co[NCOMP];
  __device__ __forceinline__ float qpos(int j) const { return qp[j]; }
  __device__ __forceinline__ float qvel(int j) const { return qv[j]; }
  __device__ __forceinline__ float act(int) const { return 0.0f; }
  __device__ __forceinline__ void obs(int j, float v) {  // j is a compile-time constant after unrolling
    if (j >= 2 * Rb::NJ) co[j - 2 * Rb::NJ] = v;         // the joint blocks are copied from the staged rows
  }
  __device__ __forceinline__ void tau(int, float) {}
};

// groups per CTA the kernel is compiled for (register budget = 65536 / (128 G) per thread)
template <class Rb>
struct KernelGroups {  // measured on humanoid3d: G = 5 (96 registers, ~330 B of spills) is 40 % slower than G = 4
  static constexpr int G = 4 * 128 / DM_TE;
  // batches of at most WIDE_WANT tiles per SM run the instantiation compiled for GWIDE groups (no register cap)
  // (with the L2 prefetch of round 2 the wide shape also wins at four tiles per SM: 65536 envs 17.9 -> 17.2 us walk,
  // 18.7 -> 17.2 us spinkick + dance_b; at seven tiles per SM it loses, 30.8 vs 28.4 us)
  static constexpr int GWIDE = 2, WIDE_WANT = 4;
  // in between, three groups at 164 registers (no spills) plus ONE lookahead ring slot beat four capped groups:
  // 262144 envs 52.2 vs 55.5 us, 524288 envs 91.0 vs 94.8, 1 M envs 168.1 vs 172.7; at 4 M envs 639 vs 634 (four groups again)
  // (after the instruction cuts of round 2 the middle shape also wins beyond 80 tiles per SM: 2 M envs 313 -> 280 us = 0.91,
  // so it has no upper bound any more; DM_MID_HI restores one)
  static constexpr int GMID = 3, MID_WANT_LO = 8, MID_WANT_HI = 1 << 30, MID_NBUF = 4;
};
template <>
struct KernelGroups<Robot_unitree_g1> {  // 37 hinges: the joint blocks alone are 74 live registers when the tile is written
  static constexpr int G = 3 * 128 / DM_TE;  // measured at 1 M envs: G = 3 (168 regs) 407 us, G = 2 (254 regs, no spills) 484 us, G = 4 (128 regs) 462 us
  // two wide groups against three capped ones: 65536 envs 29.1 vs 33.2 us, 131072 envs 52.7 vs 53.2, 262144 envs 87.0 vs 91.4,
  // 1 M envs 320 vs 308 (a lookahead slot beside the two wide groups is far slower: 108 us at 262144 envs)
  // (after the packed-arithmetic changes of round 2 the two wide groups also win at 524288 envs: combined 157.3 -> 147.9 us,
  // DPEnv 138.6 -> 137.0; at 1 M envs combined 294.2 -> 290.7 but DPEnv 256.8 -> 269.2, so the switch stays below 1 M)
  static constexpr int GWIDE = 2, WIDE_WANT = 32;
  static constexpr int GMID = 0, MID_WANT_LO = 0, MID_WANT_HI = 0, MID_NBUF = 0;  // no middle shape
};

struct TileExtra {
  double* acc;       // statistics accumulators or NULL
  int n_tiles;       // full tiles in this launch
  int table_floats;  // packed clip table size (floats); staged in shared memory when TABLE_SMEM
  int nbuf;          // ring slots (>= groups per CTA); nbuf - groups tiles are always loading ahead of the groups
  int skip_compute;  // experiment: memory skeleton only (no env_step)
  int n_cta;         // CTAs that walk tiles: blockIdx.x < n_cta; CTA n_cta is the exchange CTA (if any)
  int prefetch_next; // L2 prefetch of a slot's next tile (DM_PREFETCH_NEXT=0 switches it off)
  int no_prefetch;   // experiment (DM_NO_PREFETCH): no L2 prefetch of the first tile in front of griddepcontrol.wait
  PeerParams peer;   // peer.enable: the launch has one CTA more than tiles need, the statistics exchange CTA (dm_peer.cuh)
};

__device__ __forceinline__ void group_sync(int group) {  // named barrier 1 + group, one group of threads
  asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "n"(DM_TE) : "memory");
}

// TABLE: where a step finds its mocap row.  0 = the packed table in global memory (L2), 1 = the whole table staged once
// per CTA in shared memory, 2 = the table stays in global memory but every environment's OWN row of the group's next tile
// is gathered into a per-group stage by one small bulk copy per thread (272 B), issued a tile ahead as soon as the
// tile's clip / frame scalars have arrived -- for tables that do not fit beside the ring (G1 combined: 130 KB) this
// takes the L2 gather latency out of the step (ncu: 12 % of the G1 kernel's stall samples sat on those loads).
template <int KIND, class Rb, bool TERMS, bool STATS, int TABLE, int G>
__global__ void __launch_bounds__(DM_TE * G, 1) step_tile_kernel(const __grid_constant__ StepParams p, const TileExtra x) {
  using OL = ObsLayout<KIND, Rb>;
  using TC = TileCfg<KIND, Rb>;
  using RL = RowLayout<Rb>;
  constexpr bool TABLE_SMEM = TABLE == 1, ROW_STAGE = TABLE == 2;
  constexpr int TE = TC::TE, NJ = Rb::NJ, DIM = OL::DIM, NCOMP = TC::NCOMP, NCOLIT = TC::NCOLIT;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);      // [0..MAXBUF) ring-slot barriers, [MAXBUF] table barrier
  float* s_motor = reinterpret_cast<float*>(smem_raw + 128);   // [3][NUP]
  // issued[slot] = index j of the tile whose load was issued into the slot last.  Needed when the ring has more slots than
  // groups (the middle shape): consecutive tenants of a slot are then consumed by DIFFERENT groups, and a consumer that
  // reached its parity wait before the previous tenant's load had even completed would be let through at once
  // (mbarrier.try_wait.parity takes the parity of the immediately PRECEDING phase for "complete"), work on stale data and
  // corrupt the barrier's transaction count: an `unspecified launch failure` after some thousands of tiles, first seen as
  // a failure after 21-23 consecutive steps of 1 M environments in an experimental build and reproducible with the
  // shipped arithmetic at 2-4 M environments.  A consumer therefore first waits until its tile's load has been ISSUED --
  // which the previous consumer of the slot does after it has finished with the slot -- and only then on the parity.
  volatile int* s_issued = reinterpret_cast<volatile int*>(smem_raw + 128 + sizeof(float) * 3 * TC::NUP);
  float* s_table = reinterpret_cast<float*>(smem_raw + TC::FIXED_BYTES);
  const int n_groups = blockDim.x / TE, nbuf = x.nbuf;
  const int tid = threadIdx.x, group = tid / TE, gt = tid % TE, lane = tid & 31, warp = gt >> 5;
  float* s_ring = s_table + (TABLE_SMEM ? x.table_floats : 0);  // nbuf slots of BUF_FLOATS
  // running sums of this group: [2 NCOLIT][TE] per-thread column sums / sums of squares, then n_done, sum ep_return,
  // sum ep_len of finished episodes, tiles processed
  double* g_acc = reinterpret_cast<double*>(s_ring + (size_t)nbuf * TC::BUF_FLOATS) + (size_t)group * (2 * NCOLIT * TE + 4);
  double* my_acc = g_acc + gt;
  double* g_done = g_acc + 2 * NCOLIT * TE;
  // ROW_STAGE: [TE][RL::STRIDE] floats per group behind the running sums; its mbarrier is bars[MAXBUF + 1 + group] (<= 3 groups)
  float* s_row = reinterpret_cast<float*>(reinterpret_cast<char*>(s_ring + (size_t)nbuf * TC::BUF_FLOATS) +
                                          (STATS ? (size_t)n_groups * TC::ACC_BYTES : 0)) + (size_t)group * TE * RL::STRIDE;
  uint64_t* rbar = bars + TC::MAXBUF + 1 + (group < 3 ? group : 2);
  // this CTA's tiles are blockIdx.x + j * ncta, j = 0 .. nt-1; tile j sits in ring slot j % nbuf and is
  // processed by group j % n_groups
  const int ncta = x.n_cta;  // CTAs that walk tiles
  const int nt = (int)blockIdx.x < x.n_tiles ? (x.n_tiles - 1 - (int)blockIdx.x) / ncta + 1 : 0;
  auto issue_load = [&](int j) {  // one thread
    const int slot = j % nbuf;
    const long long f0 = ((long long)blockIdx.x + (long long)j * ncta) * TE;
    float* dst = s_ring + (size_t)slot * TC::BUF_FLOATS;
    const bool act = TC::ACT_FLOATS != 0 && p.torque != nullptr;
    mbar_expect_tx(bars + slot, sizeof(float) * (TE * (Rb::NQ + Rb::NV) + (act ? TC::ACT_FLOATS : 0)));
    tma_load_1d(dst, p.qpos + f0 * Rb::NQ, sizeof(float) * TE * Rb::NQ, bars + slot);
    tma_load_1d(dst + TE * Rb::NQ, p.qvel + f0 * Rb::NV, sizeof(float) * TE * Rb::NV, bars + slot);
    if (act) tma_load_1d(dst + TE * TC::ROW, p.action + f0 * Rb::NA, sizeof(float) * TC::ACT_FLOATS, bars + slot);
    __threadfence_block();  // the arrive above is ordered before the flag
    s_issued[slot] = j;
  };

  if (tid == 0) {
    for (int g = 0; g < 16; ++g) mbar_init(bars + g, 1);
    mbar_fence_init();
  }
  if (tid < 16) s_issued[tid] = -1;
  if (STATS) {
#pragma unroll
    for (int k = 0; k < 2 * NCOLIT; ++k) my_acc[k * TE] = 0.0;
    if (gt < 4) g_done[gt] = 0.0;
  }
  __syncthreads();
  if (STATS && x.peer.enable && (int)blockIdx.x == ncta) {
    // The statistics exchange CTA (auto mode of the collective): once the previous step is complete it collects the
    // exchange posted a step ago and posts the sums accumulated since, beside the tile CTAs' work on this step.
    if (x.peer.enable == 2) return;  // debug (DM_PEER_DEBUG): an idle extra CTA
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (x.peer.enable == 3) return;  // debug: ... that only waits for the previous kernel
    peer_fused_cta(x.peer);
    return;
  }
  // Everything up to here, the clip table and the motor constants touch nothing an earlier launch writes.  Under
  // programmatic dependent launch this prologue (and the launch latency in front of it) overlaps the tail of the
  // previous kernel in the stream; griddepcontrol.wait then blocks until that kernel has completed and its writes
  // (the in-out state of the previous step, the statistics) are visible.  Without the launch attribute both
  // instructions are no-ops.
  if (TABLE_SMEM && tid == 0) {
    mbar_expect_tx(bars + TC::MAXBUF, sizeof(float) * x.table_floats);
    tma_load_1d(s_table, p.cs.rows, sizeof(float) * x.table_floats, bars + TC::MAXBUF);
  }
  const float* rows = TABLE_SMEM ? s_table : p.cs.rows;

  // motor constants (gear, ctrlrange) of every actuator, looked up by the flat torque pass
  for (int k = tid; k < TC::NUP; k += blockDim.x) {
    float g = 0.0f, lo = 0.0f, hi = 0.0f;
#pragma unroll
    for (int kk = 0; kk < Rb::NU; ++kk)
      if (kk == k) { g = Rb::gear(kk); lo = Rb::ctrl_lo(kk); hi = Rb::ctrl_hi(kk); }
    s_motor[k] = g;
    s_motor[TC::NUP + k] = lo;
    s_motor[2 * TC::NUP + k] = hi;
  }
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // While the previous kernel of the stream finishes: pull the group's first tile towards L2.  A prefetch is a hint and L2
  // is the point of coherence, so this is safe even when that kernel is still writing these very rows.
  if (gt == 0 && group < nt && !x.no_prefetch) {
    const long long f0 = ((long long)blockIdx.x + (long long)group * ncta) * TE;
    l2_prefetch(p.qpos + f0 * Rb::NQ, sizeof(float) * TE * Rb::NQ);
    l2_prefetch(p.qvel + f0 * Rb::NV, sizeof(float) * TE * Rb::NV);
    if (p.torque) l2_prefetch(p.action + f0 * Rb::NA, sizeof(float) * TE * Rb::NA);
  }
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (gt == 0) {  // fill the ring: every leader its own first tile, leader 0 the slots beyond the groups
    if (group < nt) issue_load(group);
    if (group == 0)
      for (int j = n_groups; j < nbuf && j < nt; ++j) issue_load(j);
  }
  __syncthreads();

  bool table_ready = !TABLE_SMEM;
  // Software pipeline through registers: the per-env scalars and (128-bit path) the action chunks of a tile are
  // loaded one tile AHEAD, right after the previous tile's observation store is issued, so their global-load
  // latency hides behind the statistics pass and the barrier instead of stalling every warp of the group.
  // (Running the whole torque pass of the next tile there instead was measured slower: 188 vs 179 us at 1 M envs.)
  constexpr bool kVecTorque = TC::kVecTorque;
  constexpr int K4 = Rb::NA / 4, CH = TE * Rb::NA / 4, PER = kVecTorque ? (CH + TE - 1) / TE : 1;
  float4 av[PER];
  EnvState st_n{0, 0, 0};
  uint32_t contact_n = 0u;
  float ep_ret_n = 0.0f;
  auto load_ahead = [&](int t) {
    const long long f0 = (long long)t * TE;
    const int e = int(f0) + gt;
    st_n.idx = ld_stream(p.idx + e);
    st_n.ep_len = ld_stream(p.ep_len + e);
    st_n.clip = p.clip ? ld_stream(p.clip + e) : 0;
    contact_n = p.contact ? ld_stream(p.contact + e) : 0u;
    ep_ret_n = p.ep_return ? ld_stream(p.ep_return + e) : 0.0f;
    if (kVecTorque && p.torque) {
      const float4* a4 = reinterpret_cast<const float4*>(p.action + f0 * Rb::NA);
#pragma unroll
      for (int m = 0; m < PER; ++m)
        if (gt + m * TE < CH) av[m] = ld_stream4(a4 + gt + m * TE);
    }
  };
  // action -> torque of tile t: a flat, fully coalesced pass over the [TE, NA] tile (contiguous in the row-major
  // action array)
  auto torque_tile = [&](int t, const float* s_act) {
    const long long e0 = (long long)t * TE;
    if (p.torque) {
      if (kVecTorque) {  // 128-bit path: chunk c covers actuators 4 (c mod NA/4) .. + 3; av[] was loaded by load_ahead
        float4* t4 = reinterpret_cast<float4*>(p.torque + e0 * Rb::NU);
        const int r4 = gt % K4;
#pragma unroll
        for (int m = 0; m < PER; ++m) {
          const int c = gt + m * TE;
          if (c < CH) {
            int km = r4 + (m * TE) % K4;
            km -= km >= K4 ? K4 : 0;
            const int k0 = km * 4;
            const float4 g = *reinterpret_cast<const float4*>(s_motor + k0);
            const float4 lo = *reinterpret_cast<const float4*>(s_motor + TC::NUP + k0);
            const float4 hi = *reinterpret_cast<const float4*>(s_motor + 2 * TC::NUP + k0);
            float4 o;
            float c0 = av[m].x * Rb::kActScale, c1 = av[m].y * Rb::kActScale, c2 = av[m].z * Rb::kActScale, c3 = av[m].w * Rb::kActScale;
            // mju_clip with NaN passing through: min.NaN / max.NaN, one instruction each
            o.x = g.x * dm_max_nan(dm_min_nan(c0, hi.x), lo.x);
            o.y = g.y * dm_max_nan(dm_min_nan(c1, hi.y), lo.y);
            o.z = g.z * dm_max_nan(dm_min_nan(c2, hi.z), lo.z);
            o.w = g.w * dm_max_nan(dm_min_nan(c3, hi.w), lo.w);
            t4[c] = o;
          }
        }
      } else {  // scalar flat path: output element o = (row, actuator); hand actuators get zero torque
        constexpr int TOT = TE * Rb::NU, PERS = (TOT + TE - 1) / TE, HALF = (PERS + 1) / 2;
        const float* arow = s_act;  // the tile's [TE, NA] actions, staged by issue_load
        float* trow = p.torque + e0 * Rb::NU;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float v[HALF];
#pragma unroll
          for (int m = 0; m < HALF; ++m) {
            const int o = gt + (h * HALF + m) * TE, r = o / Rb::NU, k = o - r * Rb::NU;
            v[m] = (h * HALF + m < PERS && o < TOT && k < Rb::NA) ? arow[r * Rb::NA + k] : 0.0f;
          }
#pragma unroll
          for (int m = 0; m < HALF; ++m) {
            const int o = gt + (h * HALF + m) * TE, r = o / Rb::NU, k = o - r * Rb::NU;
            if (h * HALF + m < PERS && o < TOT) {
              float c = v[m] * Rb::kActScale;
              const float hi = s_motor[2 * TC::NUP + k], lo = s_motor[TC::NUP + k];
              c = dm_max_nan(dm_min_nan(c, hi), lo);
              trow[o] = k < Rb::NA ? s_motor[k] * c : 0.0f;
            }
          }
        }
      }
    }
  };
  // ROW_STAGE: every thread fetches the packed row its environment of the NEXT tile will read (st_n: loaded by load_ahead)
  auto issue_rows = [&]() {
    const ClipDesc cd = p.cs.clip[st_n.clip];
    const int frame = (KIND == KIND_SINGLE) ? st_n.idx : (st_n.idx % cd.len);
    const int row_i = cd.row0 + (cd.fixed_row >= 0 ? cd.fixed_row : frame);
    if (gt == 0) mbar_expect_tx(rbar, sizeof(float) * TE * RL::STRIDE);
    tma_load_1d(s_row + gt * RL::STRIDE, p.cs.rows + (size_t)row_i * RL::STRIDE, sizeof(float) * RL::STRIDE, rbar);
  };
  auto tile_of = [&](int j) { return (int)blockIdx.x + j * ncta; };
  if (group < nt) {
    load_ahead(tile_of(group));
    if (ROW_STAGE) issue_rows();
  }
  int done_tiles = 0;
  for (int j = group; j < nt; j += n_groups, ++done_tiles) {
    const int slot = j % nbuf;
    const uint32_t parity = (j / nbuf) & 1;
    float* buf = s_ring + (size_t)slot * TC::BUF_FLOATS;
    const float* s_qp = buf;
    const float* s_qv = buf + TE * Rb::NQ;
    float* s_rw = buf + TE * DIM;  // the reward column of the finished tile (statistics), behind the observation rows
    const long long e0 = (long long)tile_of(j) * TE;
    const int env = int(e0) + gt;
    EnvState st = st_n;
    const uint32_t contact = contact_n;
    const float ep_ret0 = ep_ret_n;
    if (kVecTorque) torque_tile(tile_of(j), nullptr);  // while the TMA load is in flight; its action chunks were loaded a tile ahead
    // ---- per-environment step out of shared memory --------------------------------------------------
    if (!table_ready) {
      mbar_wait(bars + TC::MAXBUF, 0);
      table_ready = true;
    }
    if (nbuf != n_groups) {  // see s_issued above
      while (s_issued[slot] < j) __nanosleep(32);
      __threadfence_block();
    }
    mbar_wait(bars + slot, parity);
    // the tile this slot receives next, towards L2 while this one is computed: its TMA load then finds it there (measured:
    // humanoid3d 262144 envs 50.1 -> 46.3 us, 1 M envs 169 -> 163 us; G1 combined 87.0 -> 82.0 us, 301 -> 292 us)
    if (x.prefetch_next && gt == 0 && j + nbuf < nt) {
      const long long f1 = ((long long)blockIdx.x + (long long)(j + nbuf) * ncta) * TE;
      l2_prefetch(p.qpos + f1 * Rb::NQ, sizeof(float) * TE * Rb::NQ);
      l2_prefetch(p.qvel + f1 * Rb::NV, sizeof(float) * TE * Rb::NV);
      if (p.torque) l2_prefetch(p.action + f1 * Rb::NA, sizeof(float) * TE * Rb::NA);
    }
    if (!kVecTorque) torque_tile(tile_of(j), buf + TE * TC::ROW);  // out of the staged action tile; the stores drain under the step
    TileIO<KIND, Rb> io;
    io.qp = s_qp + gt * Rb::NQ;
    io.qv = s_qv + gt * Rb::NV;
    StepOut so;
    if (!x.skip_compute) {
      if (ROW_STAGE) {
        mbar_wait(rbar, done_tiles & 1);
        env_step<KIND, Rb, TileIO<KIND, Rb>, TERMS, true>(io, p.cs, s_row + gt * RL::STRIDE, st, contact, so);
      } else {
        env_step<KIND, Rb, TileIO<KIND, Rb>, TERMS>(io, p.cs, rows, st, contact, so);
      }
    } else {
      so.reward = so.raw_reward = io.qp[2]; so.done = 0; so.reason = 0;
#pragma unroll
      for (int j = 0; j < NCOMP; ++j) io.co[j] = io.qv[j];
    }
    finish_env<KIND, Rb, TERMS, false>(p, env, st, so);
    const float ep_ret1 = ep_ret0 + so.raw_reward;
    if (p.ep_return) p.ep_return[env] = ep_ret1;
    if (STATS && so.done) {  // finished episodes are rare: shared-memory fp64 atomics
      atomicAdd(g_done + 0, 1.0);
      atomicAdd(g_done + 1, double(ep_ret1));
      atomicAdd(g_done + 2, double(st.ep_len));
    }
    const bool guard = so.reason == R_OBS_BOUNDS;
    // ---- observation tile in place: joint blocks -> registers, sync, finished rows over the input rows ----------
    {
      float jq[NJ], jv[NJ];
#pragma unroll
      for (int j = 0; j < NJ; ++j) jq[j] = io.qp[7 + j];
#pragma unroll
      for (int j = 0; j < NJ; ++j) jv[j] = io.qv[6 + j] * 0.1f;
      group_sync(group);  // every input row of the tile is dead
      float* orow = buf + gt * DIM;
#pragma unroll
      for (int j = 0; j < NJ; ++j) orow[j] = jq[j];
#pragma unroll
      for (int j = 0; j < NJ; ++j) orow[NJ + j] = jv[j];
#pragma unroll
      for (int j = 0; j < NCOMP; ++j) orow[2 * NJ + j] = io.co[j];
      if (STATS) s_rw[gt] = so.reward;
      if (guard) {  // observation-bounds guard: the row leaves as `obs * 0.` (zeros; NaN where an entry was infinite,
                    // deepmimic_env.py:476); rare, kept off the straight-line path
#pragma unroll 1
        for (int j = 0; j < DIM; ++j) orow[j] = orow[j] * 0.0f;
      }
    }
    fence_async_smem();
    group_sync(group);
    if (gt == 0) {
      tma_store_1d(p.obs + e0 * DIM, buf, sizeof(float) * TE * DIM);
      bulk_commit();
    }
    if (j + n_groups < nt) load_ahead(tile_of(j + n_groups));
    if (STATS) {  // column sums of the finished tile: lanes walk columns, the group's warps walk rows
      float cs32[NCOLIT], css32[NCOLIT];
#pragma unroll
      for (int it = 0; it < NCOLIT; ++it) {
        cs32[it] = css32[it] = 0.0f;
        const int c = it * 32 + lane;
        if (c <= DIM) {  // column DIM is the reward
          const float* src = c < DIM ? buf + warp * DIM + c : s_rw + warp;
          const int stride = c < DIM ? TC::NW * DIM : TC::NW;
          // two rows per packed add / fma (FADD2 / FFMA2): 4 instructions per pair of values instead of 6
          unsigned long long a2 = 0ull, q2 = 0ull;
          static_assert((TE / TC::NW) % 2 == 0, "rows in pairs");
#pragma unroll 8
          for (int r = 0; r < TE / TC::NW; r += 2) {
            const unsigned long long v2 = dm_pk2(src[r * stride], src[(r + 1) * stride]);
            a2 = dm_add2(a2, v2);
            q2 = dm_fma2(v2, v2, q2);
          }
          float lo, hi;
          dm_upk2(a2, lo, hi);
          cs32[it] = lo + hi;
          dm_upk2(q2, lo, hi);
          css32[it] = lo + hi;
        }
      }
#pragma unroll
      for (int it = 0; it < NCOLIT; ++it) {
        my_acc[it * TE] += cs32[it];
        my_acc[(NCOLIT + it) * TE] += css32[it];
      }
    }
    // the group's row stage is free (every thread passed the barriers above after its env_step) and the next tile's
    // scalars have had the statistics pass to arrive
    if (ROW_STAGE && !x.skip_compute && j + n_groups < nt) issue_rows();
    group_sync(group);  // everyone is done reading the tile before a TMA load overwrites the slot
    if (gt == 0 && j + nbuf < nt) {
      bulk_wait_read0();  // ... and so is this tile's observation store
      issue_load(j + nbuf);
    }
  }
  if (gt == 0) bulk_wait_read0();  // shared memory stays valid until the last store has read it

  if (STATS && x.acc) {
    // CTA-level reduction of the shared-memory running sums, then ONE set of fp64 atomics per CTA
    if (gt == 0) g_done[3] = double(done_tiles);
    __syncthreads();
    const double* acc0 = reinterpret_cast<const double*>(s_ring + (size_t)nbuf * TC::BUF_FLOATS);
    constexpr int GS = 2 * NCOLIT * TE + 4;
    auto colsum = [&](int c, int sq) {  // column c = it * 32 + lane, over every group and warp
      const int it = c >> 5, ln = c & 31;
      double v = 0.0;
      for (int g = 0; g < n_groups; ++g)
        for (int w = 0; w < TC::NW; ++w) v += acc0[(size_t)g * GS + ((sq ? NCOLIT : 0) + it) * TE + w * 32 + ln];
      return v;
    };
    auto donesum = [&](int k) {
      double v = 0.0;
      for (int g = 0; g < n_groups; ++g) v += acc0[(size_t)g * GS + 2 * NCOLIT * TE + k];
      return v;
    };
    for (int j = tid; j < 2 * DIM + 6; j += blockDim.x) {
      double v;
      if (j == 0) v = donesum(3) * TE;           // count
      else if (j == 1) v = colsum(DIM, 0);       // sum r
      else if (j == 2) v = colsum(DIM, 1);       // sum r^2
      else if (j < 6) v = donesum(j - 3);        // n_done, sum ep_return(done), sum ep_len(done)
      else if (j < 6 + DIM) v = colsum(j - 6, 0);
      else v = colsum(j - 6 - DIM, 1);
      if (v != 0.0) atomicAdd(x.acc + j, v);
    }
  }
}

}  // namespace dm
