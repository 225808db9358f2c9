// Host-side hand-off flags of the sequence runner (csrc/sequence.cu), header-only so that the protocol can be exercised
// under ThreadSanitizer without CUDA (tests/cpp/flags_tsan_test.cpp).
#pragma once
#include <atomic>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <vector>

namespace dms_seq {

// Monotone per-frame flags.  Every flag has exactly one waiter (the thread of the tracker that needs it), so each tracker
// thread owns a mutex / condition variable pair and a setter wakes that thread alone: with one shared condition variable
// every set woke all the trackers' threads, which hurts when the ranks of a box have more tracker threads than cores.
struct Flags {
  struct Lane { std::mutex m; std::condition_variable cv; };
  std::unique_ptr<Lane[]> lanes; size_t n_lanes = 0; std::vector<char> f;
  void reset(size_t n, size_t k) {
    if (k != n_lanes) { lanes.reset(new Lane[k]); n_lanes = k; }
    f.assign(n, 0);   // no run is in progress: the workers are parked
  }
  void set(size_t i, size_t waiter) { Lane& L = lanes[waiter % n_lanes]; { std::lock_guard<std::mutex> l(L.m); f[i] = 1; } L.cv.notify_one(); }
  void wait(size_t i, size_t me, const std::atomic<int>& err) {
    Lane& L = lanes[me % n_lanes];
    std::unique_lock<std::mutex> l(L.m);
    L.cv.wait(l, [&] { return f[i] || err.load(); });
  }
  void wake() { for (size_t k = 0; k < n_lanes; ++k) { { std::lock_guard<std::mutex> l(lanes[k].m); } lanes[k].cv.notify_all(); } }
};


}  // namespace dms_seq
