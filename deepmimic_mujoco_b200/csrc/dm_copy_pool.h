// A few helper threads that share the big host-side copies of dm_step_host (pageable caller arrays <-> the pinned
// staging block).  One memcpy stream moves ~19 GB/s on the GPU box's Xeon; the three input arrays of a 4096-env step
// (1.6 MB) therefore cost ~90 us on one core, about as much as their PCIe transfer.  copy() cuts a buffer into chunks
// that the caller AND the helpers claim from an atomic counter, so a helper that wakes late costs nothing: the caller
// simply copies more of the chunks itself.  Helpers spin briefly after a job (the copies of one step come in a burst)
// and then sleep on a condition variable.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace dm {

class CopyPool {
 public:
  explicit CopyPool(int helpers) {
    for (int i = 0; i < helpers; ++i) threads_.emplace_back([this] { worker(); });
  }
  ~CopyPool() {
    {
      std::lock_guard<std::mutex> lk(m_);
      stop_ = true;
      ++job_id_;
      job_seq_.store(job_id_, std::memory_order_release);
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  CopyPool(const CopyPool&) = delete;
  CopyPool& operator=(const CopyPool&) = delete;

  struct Seg {  // one memcpy(dst, src, bytes); src == NULL zero-fills
    void* dst;
    const void* src;
    size_t bytes;
  };
  static constexpr int kMaxSegs = 16;

  // memcpy(dst, src, bytes), shared with the helpers when it is big enough to be worth a wake-up
  void copy(void* dst, const void* src, size_t bytes) {
    const Seg s{dst, src, bytes};
    copy_many(&s, 1);
  }

  // several copies as ONE job (the arrays of one chunk of a host step): a single wake-up and a single completion wait
  // for all of them, pieces of every segment claimed from the same counter
  void copy_many(const Seg* segs, int n) {
    Job j{};
    size_t total = 0;
    for (int i = 0; i < n && j.n < kMaxSegs; ++i) {
      if (!segs[i].bytes || !segs[i].dst) continue;
      j.seg[j.n] = segs[i];
      j.first[j.n] = j.chunks;
      j.chunks += (segs[i].bytes + kChunk - 1) / kChunk;
      total += segs[i].bytes;
      ++j.n;
    }
    j.first[j.n] = j.chunks;
    if (threads_.empty() || total < kMinParallel) {
      for (int i = 0; i < j.n; ++i) one(j.seg[i].dst, j.seg[i].src, j.seg[i].bytes);
      return;
    }
    {
      // a helper that woke late may still hold a snapshot of the previous job (and would claim chunk numbers of this
      // one with it): publish only when none does.  Helpers take their snapshot under this mutex.
      std::unique_lock<std::mutex> lk(m_);
      idle_.wait(lk, [this] { return active_ == 0; });
      job_ = j;
      next_.store(0, std::memory_order_relaxed);
      done_.store(0, std::memory_order_relaxed);
      ++job_id_;
      job_seq_.store(job_id_, std::memory_order_release);
    }
    cv_.notify_all();
    run(j);
    while (done_.load(std::memory_order_acquire) < j.chunks) cpu_relax();  // chunks claimed by helpers still in flight
  }

 private:
  struct Job {
    Seg seg[kMaxSegs];
    size_t first[kMaxSegs + 1];  // first piece number of every segment
    size_t chunks;
    int n;
  };
  static constexpr size_t kChunk = 32 * 1024, kMinParallel = 128 * 1024;
  static constexpr int kSpinUs = 400;  // a host step copies in at its start and (sometimes) out at its end, ~100-300 us apart; helpers that fell asleep in between wake too late to help
  static void one(void* dst, const void* src, size_t bytes) {
    if (src) memcpy(dst, src, bytes);
    else memset(dst, 0, bytes);
  }
  static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
  }
  void run(const Job& j) {
    for (;;) {
      const size_t c = next_.fetch_add(1, std::memory_order_relaxed);
      if (c >= j.chunks) return;
      int k = 0;
      while (c >= j.first[k + 1]) ++k;
      const Seg& g = j.seg[k];
      const size_t o = (c - j.first[k]) * kChunk, len = o + kChunk <= g.bytes ? kChunk : g.bytes - o;
      one((char*)g.dst + o, g.src ? (const char*)g.src + o : nullptr, len);
      done_.fetch_add(1, std::memory_order_release);
    }
  }
  void worker() {
    unsigned long long seen = 0;
    for (;;) {
      // spin for the next job of the burst, then sleep
      const auto t0 = std::chrono::steady_clock::now();
      while (job_seq_.load(std::memory_order_acquire) == seen) {
        cpu_relax();
        if (std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(kSpinUs)) break;
      }
      Job j;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return job_id_ != seen; });
        if (stop_) return;
        seen = job_id_;
        j = job_;
        ++active_;
      }
      run(j);
      {
        std::lock_guard<std::mutex> lk(m_);
        if (--active_ == 0) idle_.notify_one();
      }
    }
  }

  std::vector<std::thread> threads_;
  std::mutex m_;
  std::condition_variable cv_, idle_;
  Job job_{};
  unsigned long long job_id_ = 0;
  std::atomic<unsigned long long> job_seq_{0};
  std::atomic<size_t> next_{0}, done_{0};
  int active_ = 0;
  bool stop_ = false;
};

}  // namespace dm
