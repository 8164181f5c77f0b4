// ThreadSanitizer driver for the host-side state of the native runtime that several threads touch without a GPU in the
// loop: the all-reduce dispatcher table (mutex-guarded, replaced at run time) and the per-thread error slot behind
// dmlp_last_error. Built and run by scripts/tsan_host_runtime.sh (host code of collectives.cu / elementwise.cu compiled
// with -fsanitize=thread); needs no CUDA device. The gRPC servers call these from up to 16 worker threads.
#include <atomic>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/dmlp_b200.h"

int main() {
  std::atomic<int> bad{0};
  std::atomic<bool> stop{false};
  std::vector<std::thread> th;
  // readers: what algo=auto would run, from many threads
  for (int t = 0; t < 8; ++t)
    th.emplace_back([&, t] {
      unsigned long long sum = 0;
      while (!stop.load(std::memory_order_relaxed)) {
        for (int w = 2; w <= 8; ++w)
          for (size_t b = 1024; b <= (1ull << 30); b <<= 3) {
            const int a = dmlp_allreduce_pick_algo_ex(w, b, t & 1);
            if (a != DMLP_ALGO_ONESHOT && a != DMLP_ALGO_TWOSHOT && a != DMLP_ALGO_NVLS) bad.fetch_add(1);
            sum += static_cast<unsigned long long>(a);
          }
      }
      if (sum == 0) bad.fetch_add(1);
    });
  // writers: replace and restore rows; malformed specs must fail and leave a message in THIS thread's error slot
  for (int t = 0; t < 2; ++t)
    th.emplace_back([&, t] {
      for (int i = 0; i < 2000; ++i) {
        if (dmlp_allreduce_set_table(t ? "8:256K:32M,2:1M:never" : "4:2M:128M,5:64K:1G") != DMLP_OK) bad.fetch_add(1);
        if (dmlp_allreduce_set_table(nullptr) != DMLP_OK) bad.fetch_add(1);
        if (dmlp_allreduce_set_table(t ? "8:256K" : "x:1:2") == DMLP_OK) bad.fetch_add(1);
        if (std::strstr(dmlp_last_error(), t ? "8:256K" : "x:1:2") == nullptr) bad.fetch_add(1);
      }
    });
  for (size_t i = 8; i < th.size(); ++i) th[i].join();
  stop.store(true);
  for (size_t i = 0; i < 8; ++i) th[i].join();
  dmlp_allreduce_set_table(nullptr);
  const bool restored = dmlp_allreduce_pick_algo_ex(8, 512u << 10, 0) == DMLP_ALGO_ONESHOT &&
                        dmlp_allreduce_pick_algo_ex(8, 1u << 20, 0) == DMLP_ALGO_TWOSHOT;
  std::printf("tsan_host: %d inconsistencies, table restored: %s\n", bad.load(), restored ? "yes" : "no");
  return (bad.load() == 0 && restored) ? 0 : 1;
}
