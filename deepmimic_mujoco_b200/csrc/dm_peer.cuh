// K7: the statistics collective of the path as ONE device-side exchange over NVLink peer memory.
//
// The reference has no collective (SubprocVecEnv pipes only, src/sb3_ppo.py:275-278); the reward / observation-normaliser
// statistics are new functionality.  Round 1 all-reduced the RUNNING accumulators in place with a host-enqueued
// ncclAllReduce: wrong (every further collective re-multiplied what was already summed) and, at 4096 environments per
// GPU, 3-5x the cost of the step it followed.  Here:
//   * the step kernels keep adding into the rank's accumulators `acc`;
//   * an exchange SNAPSHOTS-AND-ZEROES `acc` (atomicExch per entry: a step kernel running beside it adds with atomicAdd,
//     so every contribution lands either in this snapshot or in the next one and none is lost), writes the snapshot into
//     line set [parity][rank] of EVERY rank's communication buffer (plain stores to peer-mapped memory, i.e. NVLink
//     writes), and adds the lines of every rank -- summed in rank order, so every rank gets the same bits -- to the
//     running global `total`.  Only snapshots are ever exchanged; `total` is never re-reduced.
//   * Every value travels with its sequence number: one 16-byte line {low word, tag, high word, tag} per double, each
//     8-byte half atomic on its own (the scheme of NCCL's LL protocol).  A sender needs no release fence and no separate
//     flag -- the remote stores are fire-and-forget -- and the receiver validates each value by its two tags.
//   * No host involvement, no NCCL launch, graph-capturable (the sequence number lives in device memory).
//   * dm_stats_allreduce launches the COMPLETE exchange (send, then wait for everyone's lines of the same number).
//   * Auto mode (dm_stats_set_auto) uses the LAGGED exchange: take in the lines posted one exchange ago -- they have had
//     a whole step to arrive -- and post the sums accumulated since; nothing waits for the exchange it posts.  The
//     totals trail by one exchange while the loop runs; one dm_stats_allreduce at the end makes them exact.  It runs
//     either as its own one-CTA kernel on a side stream with NO dependency on the step's stream in either direction
//     (small launches: free SMs exist, and a kernel that has written to peer memory completes ~1.2 us later, which the
//     step's stream must not wait for), or as one extra CTA of the step kernel (launches that fill the GPU, where a
//     second kernel would have to wait for a free SM and the extra CTA is free).
//     Measured at 2 x 4096 environments, an exchange every step (profiles/r02_collective.md): N = 1 8.6 us/step;
//     ncclAllReduce of the snapshot 37; complete exchange kernel between the steps 17-20; the same behind an event on a
//     side stream 11.8; extra CTA with flags + release fences 10.0-10.2; extra CTA, tagged lines 9.9-10.0; a PULL form
//     (publish locally, read the peers) 11.8; lagged kernel on the side stream, no events **9.1**.
// Two line sets (parity of the sequence number) make the exchange safe without a barrier: a rank sends exchange seq + 1
// into set (seq + 1) & 1 only after it has received exchange seq from everyone, and a peer sent its seq only after it
// had received seq - 1, the previous tenant of that set.
// A rank that does not see its peers within kPeerTimeoutNs gives up, sets the error word and adds only what arrived: a
// missing rank must never hang the GPU.
#pragma once
#include <cstdint>

namespace dm {

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerStride = 256;  // values per line set (n_stats <= 6 + 2 * 125)
constexpr unsigned long long kPeerTimeoutNs = 2000000000ull;

struct PeerComm {  // one per rank, in that rank's device memory, mapped by every peer (cudaIpc)
  uint4 line[2][kPeerMaxWorld][kPeerStride];  // {lo, tag, hi, tag} per value; set [parity][r] is written by rank r
};

struct PeerState {  // device memory, one per handle
  unsigned long long seq;  // sequence number of the last exchange posted
  unsigned int pending;    // 1: exchange `seq` has been posted but not yet received (lagged form)
  unsigned int err;        // 1: a peer did not arrive in time
};

struct PeerParams {
  double* acc;                 // the rank's running accumulators (zeroed by a send)
  double* snap;                // the rank's last snapshot (kept for inspection / the NCCL fallback)
  double* total;               // global running sums
  PeerState* st;
  PeerComm* comm[kPeerMaxWorld];  // comm[r] = rank r's buffer as mapped in THIS process (comm[rank] = own)
  int rank, world, n;
  int enable;                  // step kernel: this launch carries an exchange CTA
};

__device__ __forceinline__ unsigned long long peer_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void ll_store(uint4* p, double v, uint32_t tag) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((uint32_t)b), "r"(tag), "r"((uint32_t)(b >> 32)), "r"(tag)
               : "memory");
}
__device__ __forceinline__ uint4 ll_load(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// snapshot-and-zero, then the snapshot towards every rank (the own lines too: the receiver treats all ranks alike).
// One rank: the snapshot goes straight to the totals.
__device__ __forceinline__ void ll_send(const PeerParams& p, unsigned long long seq) {
  const int par = int(seq & 1);
  const uint32_t tag = (uint32_t)seq;
  for (int j = threadIdx.x; j < p.n; j += blockDim.x) {
    const double v = __longlong_as_double((long long)atomicExch(reinterpret_cast<unsigned long long*>(p.acc + j), 0ull));
    p.snap[j] = v;
    if (p.world > 1) {
      for (int r = 0; r < p.world; ++r) ll_store(&p.comm[r]->line[par][p.rank][j], v, tag);
    } else {
      p.total[j] += v;
    }
  }
}
// wait for exchange `seq` value by value and add it, in rank order, to the totals
__device__ __forceinline__ void ll_recv(const PeerParams& p, unsigned long long seq) {
  if (p.world <= 1) return;
  const int par = int(seq & 1);
  const uint32_t tag = (uint32_t)seq;
  bool late = false;
  for (int j = threadIdx.x; j < p.n; j += blockDim.x) {
    uint4 v[kPeerMaxWorld];
    const unsigned long long t0 = peer_now();
    for (;;) {
      bool ok = true;
#pragma unroll
      for (int r = 0; r < kPeerMaxWorld; ++r)
        if (r < p.world) {
          v[r] = ll_load(&p.comm[p.rank]->line[par][r][j]);
          ok = ok && v[r].y == tag && v[r].w == tag;
        }
      if (ok) break;
      if (peer_now() - t0 > kPeerTimeoutNs) {
        late = true;
        break;
      }
      __nanosleep(64);
    }
    double s = 0.0;
#pragma unroll
    for (int r = 0; r < kPeerMaxWorld; ++r)
      if (r < p.world && v[r].y == tag && v[r].w == tag)
        s += __longlong_as_double((long long)(((unsigned long long)v[r].z << 32) | v[r].x));
    p.total[j] += s;
  }
  if (late) p.st->err = 1u;
  __syncthreads();  // every line of exchange `seq` has been read before this rank sends again
}

// One exchange by one CTA.  complete = false: the LAGGED form (receive what was posted one exchange ago, post the sums
// accumulated since); complete = true: additionally wait for the exchange just posted, which makes the totals exact.
__device__ __forceinline__ void peer_exchange(const PeerParams& p, bool complete) {
  __shared__ unsigned long long s_seq;
  __shared__ unsigned int s_pending;
  if (threadIdx.x == 0) {
    s_seq = p.st->seq;
    s_pending = p.st->pending;
  }
  __syncthreads();
  const unsigned long long seq = s_seq;
  if (s_pending) ll_recv(p, seq);
  ll_send(p, seq + 1);
  if (complete) ll_recv(p, seq + 1);
  if (threadIdx.x == 0) {
    p.st->seq = seq + 1;
    p.st->pending = (!complete && p.world > 1) ? 1u : 0u;
  }
}
// the exchange CTA of the step kernel (dm_tile_kernel.cuh)
__device__ __forceinline__ void peer_fused_cta(const PeerParams& p) { peer_exchange(p, false); }
// the lagged exchange as its own kernel, for the side stream of the auto mode
static __global__ void __launch_bounds__(kPeerStride) stats_exchange_lagged_kernel(const PeerParams p) { peer_exchange(p, false); }
// the explicit exchange (dm_stats_allreduce): receive what the auto mode left pending, then a complete exchange
static __global__ void __launch_bounds__(kPeerStride) stats_exchange_kernel(const PeerParams p) { peer_exchange(p, true); }

// pieces of the checked fallback (NCCL all-reduce of the SNAPSHOT between them): acc -> snap (zeroing acc); total += snap
static __global__ void stats_snapshot_kernel(double* acc, double* snap, int n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) {
    snap[j] = acc[j];
    acc[j] = 0.0;
  }
}
static __global__ void stats_commit_kernel(const double* snap, double* total, int n) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) total[j] += snap[j];
}

}  // namespace dm
