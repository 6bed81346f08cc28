// Open-loop load generator at the C ABI (measurement aid; no reference counterpart - BASELINE.json configs[4] asks
// for Poisson arrivals at 50 % / 80 % of peak).  Python cannot offer eight GPUs their load one
// image at a time (one interpreter tops out near 330 k calls/s under the GIL, tools/openloop_bench.py), so the
// arrivals are generated here, by native threads, through the same public entry points a host shell uses:
// ai4e_pool_submit_many (one call per wake-up with every arrival of this feeder that is due, tag = arrival number,
// notify) and ai4e_pool_poll.  Latency of an arrival = time its result was announced - its SCHEDULED time, so time
// spent waiting for a feeder or for the micro-batch to close is counted.
#include "../../include/ai4e_b200.h"
#include "common.cuh"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <random>
#include <thread>
#include <vector>

extern "C" int ai4e_pool_openloop(ai4e_pool* p, const uint8_t* h_images, int n_images, int h, int w, int c,
                                  double rate_per_s, double seconds, int feeders, uint64_t seed, float* out_latency_ms,
                                  int64_t capacity, int64_t* out_offered, int64_t* out_completed, int64_t* out_failed,
                                  double* out_span_s, double* out_feeder_lateness_ms) {
  AI4E_REQUIRE(p && h_images && out_latency_ms && out_offered && out_completed && out_failed && out_span_s && out_feeder_lateness_ms,
               "pool_openloop: null pointer");
  AI4E_REQUIRE(n_images >= 1 && h > 0 && w > 0 && c > 0 && rate_per_s > 0 && seconds > 0 && feeders >= 1 && feeders <= 256,
               "pool_openloop: bad arguments");
  const int64_t n = std::min<int64_t>((int64_t)(rate_per_s * seconds), capacity);
  AI4E_REQUIRE(n >= 1, "pool_openloop: no arrivals (rate %.1f/s for %.3f s, capacity %lld)", rate_per_s, seconds, (long long)capacity);
  using Clock = std::chrono::steady_clock;
  std::vector<double> arrival((size_t)n);
  {
    std::mt19937_64 rng(seed);
    std::exponential_distribution<double> gap(rate_per_s);
    double t = 0;
    for (int64_t i = 0; i < n; ++i) { t += gap(rng); arrival[(size_t)i] = t; }
  }
  std::vector<double> done((size_t)n, -1.0);
  std::vector<int32_t> status((size_t)n, 0);
  std::atomic<int64_t> n_done{0}, n_failed_submit{0};
  std::atomic<bool> stop{false};
  std::vector<double> late((size_t)feeders, 0.0);
  const size_t img_bytes = (size_t)h * w * c;
  const auto t0 = Clock::now();
  auto now_s = [&] { return std::chrono::duration<double>(Clock::now() - t0).count(); };

  std::thread reaper([&] {
    constexpr int kCap = 4096;
    std::vector<uint64_t> tickets(kCap), tags(kCap);
    std::vector<int32_t> st(kCap);
    while (!stop.load() && n_done.load() + n_failed_submit.load() < n) {
      int got = 0;
      if (ai4e_pool_poll(p, kCap, 20000, tickets.data(), tags.data(), st.data(), nullptr, nullptr, &got) != AI4E_OK) break;
      if (got <= 0) continue;
      const double t = now_s();
      for (int i = 0; i < got; ++i) {
        if (tags[i] < (uint64_t)n) { done[tags[i]] = t; status[tags[i]] = st[i]; }
      }
      n_done.fetch_add(got);
    }
  });

  std::vector<std::thread> th;
  for (int f = 0; f < feeders; ++f) {
    th.emplace_back([&, f] {
      constexpr int kGroup = 64;
      const uint8_t* ptrs[kGroup];
      uint64_t tags[kGroup], tickets[kGroup];
      int64_t next = f;  // this feeder owns arrivals f, f + feeders, ...
      while (next < n && !stop.load()) {
        const double t = now_s();
        if (arrival[(size_t)next] > t) {
          const double gap = arrival[(size_t)next] - t;
          if (gap > 100e-6) std::this_thread::sleep_for(std::chrono::duration<double>(std::min(gap * 0.7, 500e-6)));
          continue;
        }
        late[(size_t)f] = std::max(late[(size_t)f], t - arrival[(size_t)next]);
        int k = 0;
        while (k < kGroup && next < n && arrival[(size_t)next] <= t) {
          ptrs[k] = h_images + (size_t)(next % n_images) * img_bytes;
          tags[k] = (uint64_t)next;
          ++k;
          next += feeders;
        }
        if (ai4e_pool_submit_many(p, ptrs, k, h, w, c, 0, tags, 1, tickets) != AI4E_OK) {
          n_failed_submit.fetch_add(k);
          for (int i = 0; i < k; ++i) status[tags[i]] = AI4E_ERR_STATE;
        }
      }
    });
  }
  for (auto& t : th) t.join();
  const auto deadline = Clock::now() + std::chrono::seconds(30);
  while (n_done.load() + n_failed_submit.load() < n && Clock::now() < deadline) std::this_thread::sleep_for(std::chrono::milliseconds(2));
  stop.store(true);
  reaper.join();

  int64_t completed = 0, failed = n_failed_submit.load();
  double last = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (done[(size_t)i] < 0) { out_latency_ms[i] = -1.0f; continue; }
    if (status[(size_t)i] != AI4E_OK) { ++failed; out_latency_ms[i] = -1.0f; continue; }
    out_latency_ms[i] = (float)((done[(size_t)i] - arrival[(size_t)i]) * 1e3);
    last = std::max(last, done[(size_t)i]);
    ++completed;
  }
  *out_offered = n;
  *out_completed = completed;
  *out_failed = failed;
  *out_span_s = last;
  *out_feeder_lateness_ms = *std::max_element(late.begin(), late.end()) * 1e3;
  return AI4E_OK;
}
