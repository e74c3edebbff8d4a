// K7: the statistics collective of the path as ONE device-side exchange over NVLink peer memory.
//
// The reference has no collective (SubprocVecEnv pipes only, src/sb3_ppo.py:275-278); the reward / observation-normaliser
// statistics are new functionality.  Round 1 all-reduced the RUNNING accumulators in place with a host-enqueued
// ncclAllReduce: wrong (every further collective re-multiplied what was already summed) and, at 4096 environments per
// GPU, 3-5x the cost of the step it followed.  Here:
//   * the step kernels keep adding into the rank's accumulators `acc`;
//   * dm_stats_allreduce launches this one-CTA kernel on the step's stream: it SNAPSHOTS-AND-ZEROES `acc`, writes the
//     snapshot into slot [parity][rank] of EVERY rank's communication buffer (plain stores to peer-mapped memory, i.e.
//     NVLink writes), fences, raises flag [parity][rank] = sequence number on every rank, waits until all of its own flags
//     show the sequence number, and adds the slots -- summed in rank order, so every rank gets the same bits -- to the
//     running global `total`.  Only snapshots are ever exchanged; `total` is never re-reduced.
//   * No host involvement, no NCCL launch, graph-capturable (the sequence number lives in device memory).
//   * Auto mode (dm_stats_set_auto): the exchange rides INSIDE the step kernel as one extra CTA (peer_fused_cta): at the
//     start of step k + 1's kernel it collects the exchange posted a step earlier and posts the sums of the steps up to k;
//     no extra launch, no event, no exposed NVLink latency.  Measured alternatives (profiles/r02_collective.md): a
//     separate exchange kernel between the steps 19.6 us/step at 2 x 4096 environments, the same kernel on a side stream
//     behind an event 11.8 us (the event record costs the programmatic-launch overlap), ncclAllReduce of the snapshot 37 us.
// Two slot sets (parity of the sequence number) make the exchange safe without a second barrier: a rank can start call
// k + 1 (writing set (k+1) & 1) while a slower rank still reads set k & 1, and nobody can start call k + 2 before every
// rank has raised its flag of call k + 1, i.e. has finished reading set k & 1.
// A rank that does not see its peers within kPeerTimeoutNs gives up, sets the error word and adds only what arrived: a
// missing rank must never hang the GPU.
#pragma once
#include <cstdint>

namespace dm {

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerStride = 256;  // doubles per slot (n_stats <= 6 + 2 * 125)
constexpr unsigned long long kPeerTimeoutNs = 2000000000ull;

struct PeerComm {  // one per rank, in that rank's device memory, mapped by every peer (cudaIpc)
  double slots[2][kPeerMaxWorld][kPeerStride];
  unsigned long long flags[2][kPeerMaxWorld];
};

struct PeerState {  // device memory, one per handle
  unsigned long long seq;  // sequence number of the last exchange posted
  unsigned int pending;    // 1: exchange `seq` has been posted but not yet collected (fused mode)
  unsigned int err;        // 1: a peer did not arrive in time
};

struct PeerParams {
  double* acc;                 // the rank's running accumulators (zeroed by a post)
  double* snap;                // the rank's last snapshot (kept for inspection / the NCCL fallback)
  double* total;               // global running sums
  PeerState* st;
  PeerComm* comm[kPeerMaxWorld];  // comm[r] = rank r's buffer as mapped in THIS process (comm[rank] = own)
  int rank, world, n;
  int enable;                  // fused mode: this launch carries an exchange CTA
};

__device__ __forceinline__ unsigned long long peer_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void peer_store_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_load_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long peer_load_relaxed(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// The four phases of an exchange, each executed by the whole CTA.  Ordering: the flag of exchange `seq` is stored with
// release semantics after a CTA barrier that follows every thread's slot stores (cumulativity through bar.sync), and a
// reader polls the flag relaxed, fences once, and passes a CTA barrier before anyone reads the slots.
//
// (1) wait until every rank has posted exchange `seq`
__device__ __forceinline__ void peer_wait(const PeerParams& p, unsigned long long seq, int* s_late) {
  if (p.world > 1 && threadIdx.x < p.world) {
    const unsigned long long t0 = peer_now();
    const unsigned long long* f = &p.comm[p.rank]->flags[int(seq & 1)][threadIdx.x];
    while (peer_load_relaxed(f) < seq) {
      __nanosleep(64);
      if (peer_now() - t0 > kPeerTimeoutNs) {
        atomicExch(s_late, 1);
        break;
      }
    }
    __threadfence_system();
  }
  __syncthreads();
}
// (2) snapshot-and-zero the accumulators and send the snapshot towards every rank's slot of exchange `seq`.  The snapshot
// is atomic per entry: a step kernel running beside this (fused mode) adds with atomicAdd, so every contribution lands
// either in this snapshot or in the next one and none is lost.  One rank: the snapshot goes straight to the totals.
__device__ __forceinline__ void peer_send(const PeerParams& p, unsigned long long seq) {
  const int par = int(seq & 1);
  for (int j = threadIdx.x; j < p.n; j += blockDim.x) {
    const double v = __longlong_as_double((long long)atomicExch(reinterpret_cast<unsigned long long*>(p.acc + j), 0ull));
    p.snap[j] = v;
    if (p.world > 1) {
      for (int r = 0; r < p.world; ++r) p.comm[r]->slots[par][p.rank][j] = v;
    } else {
      p.total[j] += v;
    }
  }
}
// (3) add the slots of exchange `seq` (all present: peer_wait) to the global totals, in rank order: the same bits everywhere
__device__ __forceinline__ void peer_sum(const PeerParams& p, unsigned long long seq, const int* s_late) {
  if (p.world <= 1) return;
  const int par = int(seq & 1);
  for (int j = threadIdx.x; j < p.n; j += blockDim.x) {
    double s = 0.0;
    for (int r = 0; r < p.world; ++r) s += __ldcv(&p.comm[p.rank]->slots[par][r][j]);  // L1 may hold the lines of exchange seq - 2
    p.total[j] += s;
  }
  if (threadIdx.x == 0 && *s_late) p.st->err = 1u;
}
// (4) tell every rank that this rank's slot of exchange `seq` is complete
__device__ __forceinline__ void peer_raise(const PeerParams& p, unsigned long long seq) {
  __syncthreads();
  if (p.world > 1 && threadIdx.x < p.world) peer_store_release(&p.comm[threadIdx.x]->flags[int(seq & 1)][p.rank], seq);
}

// The exchange CTA of the FUSED mode (one extra CTA of the step kernel, after griddepcontrol.wait: the previous step is
// complete): take in the exchange posted one step ago -- its data has had a whole step to arrive, so the wait is a
// formality -- and post the sums accumulated since.  The summation of the old exchange runs while the stores of the new
// one cross NVLink, and nothing here waits for the exchange it posts: its latency lies behind the step's own compute.
// The totals therefore trail by one exchange; dm_stats_allreduce (the explicit kernel below) takes in what is pending
// and exchanges the remainder, which makes them exact.
// Slot-set safety: set (seq+1) & 1 was last used by exchange seq - 1.  A peer raised its flag of exchange `seq` only
// after it had summed seq - 1 (program order here), so having waited for `seq` this rank may overwrite that set.
__device__ __forceinline__ void peer_fused_cta(const PeerParams& p) {
  __shared__ unsigned long long s_seq;
  __shared__ unsigned int s_pending;
  __shared__ int s_late;
  if (threadIdx.x == 0) {
    s_seq = p.st->seq;
    s_pending = p.st->pending;
    s_late = 0;
  }
  __syncthreads();
  const unsigned long long seq = s_seq;
  const bool pending = s_pending != 0;
  if (pending) peer_wait(p, seq, &s_late);
  peer_send(p, seq + 1);
  if (pending) peer_sum(p, seq, &s_late);
  peer_raise(p, seq + 1);
  if (threadIdx.x == 0) {
    p.st->seq = seq + 1;
    p.st->pending = p.world > 1 ? 1u : 0u;
  }
}

// The explicit exchange (dm_stats_allreduce): collect what the fused mode left pending, then a complete exchange.
static __global__ void __launch_bounds__(kPeerStride) stats_exchange_kernel(const PeerParams p) {
  __shared__ unsigned long long s_seq;
  __shared__ unsigned int s_pending;
  __shared__ int s_late;
  if (threadIdx.x == 0) {
    s_seq = p.st->seq;
    s_pending = p.st->pending;
    s_late = 0;
  }
  __syncthreads();
  const unsigned long long seq = s_seq;
  const bool pending = s_pending != 0;
  if (pending) peer_wait(p, seq, &s_late);
  peer_send(p, seq + 1);
  if (pending) peer_sum(p, seq, &s_late);
  peer_raise(p, seq + 1);
  peer_wait(p, seq + 1, &s_late);
  peer_sum(p, seq + 1, &s_late);
  if (threadIdx.x == 0) {
    p.st->seq = seq + 1;
    p.st->pending = 0u;
  }
}

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
