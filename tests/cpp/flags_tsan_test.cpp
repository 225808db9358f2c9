// The host-side hand-off protocol of the sequence runner (csrc/sequence.cu: seq_work) on plain memory, for ThreadSanitizer.
//
// K tracker threads take turns on N frames.  Frame i (thread i % K) overwrites its tracker's image slot with frame i, which it
// may do only after the image the slot held (frame i - K) has been taken by frame i - K + 1; then it takes the image of frame
// i - 1 from the slot of tracker (i - 1) % K.  In the library the accesses are device work ordered by CUDA events and the
// flags order the host calls that record / wait for those events; here the slots are ordinary ints, so a missing
// happens-before edge in the flag protocol is a data race ThreadSanitizer reports, and a wrong value is an ordering bug.
//
//   g++ -std=c++17 -O1 -g -fsanitize=thread -pthread -I densemonoslam_b200/csrc tests/cpp/flags_tsan_test.cpp -o flags_tsan_test
#include <stdio.h>
#include <stdlib.h>
#include <atomic>
#include <chrono>
#include <random>
#include <thread>
#include <vector>
#include "seq_flags.h"

int main(int argc, char** argv) {
  const int K = argc > 1 ? atoi(argv[1]) : 7, N = argc > 2 ? atoi(argv[2]) : 400, runs = argc > 3 ? atoi(argv[3]) : 5;
  dms_seq::Flags issued, taken;
  std::atomic<int> err{0};
  std::vector<int> slot((size_t)K, -1);   // frame whose image tracker k holds (plain memory on purpose)
  long long bad = 0;
  for (int run = 0; run < runs; ++run) {
    issued.reset((size_t)N + 1, (size_t)K);
    taken.reset((size_t)N + K + 1, (size_t)K);
    for (int j = 0; j < K; ++j) taken.set((size_t)j, (size_t)j);
    for (int k = 0; k < K; ++k) slot[(size_t)k] = -1;
    std::vector<long long> bad_k((size_t)K, 0);
    std::vector<std::thread> th;
    for (int k = 0; k < K; ++k)
      th.emplace_back([&, k] {
        std::minstd_rand rng((unsigned)(k * 7919 + run));
        for (int i = k; i < N; i += K) {
          if (rng() % 4 == 0) std::this_thread::sleep_for(std::chrono::microseconds(rng() % 50));   // the init calls before
          if (K > 1) taken.wait((size_t)i, (size_t)k, err);
          slot[(size_t)k] = i;                                                   // initRGB(frame i): the tracker's pyramid now holds frame i
          issued.set((size_t)i, (size_t)(i + 1) % K);
          if (i > 0 && K > 1) {
            issued.wait((size_t)i - 1, (size_t)k, err);
            if (slot[(size_t)((i - 1) % K)] != i - 1) ++bad_k[(size_t)k];        // init_first_rgb_from(tracker of frame i - 1)
          }
          taken.set((size_t)i - 1 + K, (size_t)(i - 1 + K) % K);
          if (rng() % 3 == 0) std::this_thread::yield();                         // track_async ... collect
        }
      });
    for (auto& t : th) t.join();
    for (long long b : bad_k) bad += b;
  }
  // the error path: a waiter of a flag that never comes must return once err is set and wake() is called
  issued.reset(4, 2);
  std::thread w([&] { issued.wait(3, 1, err); });
  std::this_thread::sleep_for(std::chrono::milliseconds(20));
  err.store(1); issued.wake();
  w.join();
  printf("flags protocol: K=%d N=%d runs=%d ordering violations=%lld\n", K, N, runs, bad);
  return bad == 0 ? 0 : 1;
}
