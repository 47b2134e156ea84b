// TEST INFRASTRUCTURE ONLY (see oracle/cpu_oracle.py): multi-threaded host twins of two input
// builders, so that the CPU legs of bench.py can run the reference's pagerank_reference on the
// SAME graph as the GPU arm (R-MAT scale 24 = 268 M edges) within seconds instead of minutes:
//   oracle_rmat_edgelist_i32  the bit loop of cpp/src/generators/generate_rmat_edgelist.cu:75-107
//                             over this repository's counter-based stream (splitmix64; identical
//                             to oracle/graph_oracle.py rmat_edgelist and csrc/rmat.cu);
//   oracle_csc_from_coo_i32   un-renumbered CSC in input order, i.e. a stable counting sort by
//                             destination - what numpy's stable argsort gives in
//                             cpu_oracle.csc_from_coo (pagerank_test.cpp:264-300 builds the host
//                             graph the same way).
// tests/test_oracle.py pins both against the numpy versions.
#include <algorithm>
#include <cstdint>
#include <thread>
#include <vector>

namespace {
inline uint64_t splitmix64(uint64_t x)
{
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline float uniform01(uint64_t counter, uint64_t seed)
{
  return static_cast<float>(splitmix64(counter + seed) >> 40) * 5.9604644775390625e-8f;
}
template <typename F>
void parallel_for(int64_t n, int threads, F&& f)
{
  threads = std::max(1, threads);
  std::vector<std::thread> pool;
  const int64_t per = (n + threads - 1) / threads;
  for (int t = 0; t < threads; ++t) {
    const int64_t a = std::min<int64_t>(n, t * per), b = std::min<int64_t>(n, a + per);
    if (a < b) pool.emplace_back([=, &f] { f(t, a, b); });
  }
  for (auto& th : pool) th.join();
}
}  // namespace

extern "C" {

void oracle_rmat_edgelist_i32(int scale, int64_t num_edges, double a, double b, double c,
                              uint64_t seed, uint64_t first_edge, int32_t* src, int32_t* dst,
                              int threads)
{
  const double a_plus_b = a + b;
  const double a_norm   = (a + b) > 0.0 ? a / (a + b) : 0.0;
  const double c_norm   = (1.0 - (a + b)) > 0.0 ? c / (1.0 - (a + b)) : 0.0;
  parallel_for(num_edges, threads, [&](int, int64_t lo, int64_t hi) {
    for (int64_t i = lo; i < hi; ++i) {
      const uint64_t base = (first_edge + static_cast<uint64_t>(i)) * static_cast<uint64_t>(2 * scale);
      int32_t s = 0, d = 0;
      for (int bit = scale - 1; bit >= 0; --bit) {
        const float r0 = uniform01(base + 2 * bit, seed);
        const float r1 = uniform01(base + 2 * bit + 1, seed);
        const bool sb  = static_cast<double>(r0) > a_plus_b;
        const bool db  = static_cast<double>(r1) > (sb ? c_norm : a_norm);
        s += sb ? (int32_t{1} << bit) : 0;
        d += db ? (int32_t{1} << bit) : 0;
      }
      src[i] = s;
      dst[i] = d;
    }
  });
}

// offsets[nv + 1] (int64), indices[ne] = src in stable order of dst
void oracle_csc_from_coo_i32(const int32_t* src, const int32_t* dst, int64_t ne, int64_t nv,
                             int64_t* offsets, int32_t* indices, int threads)
{
  threads = std::max(1, threads);
  // per-thread histograms over contiguous edge ranges keep the sort stable
  std::vector<std::vector<int64_t>> hist(threads);
  const int64_t per = (ne + threads - 1) / threads;
  parallel_for(ne, threads, [&](int t, int64_t lo, int64_t hi) {
    hist[t].assign(static_cast<size_t>(nv), 0);
    for (int64_t e = lo; e < hi; ++e) ++hist[t][dst[e]];
  });
  for (auto& h : hist) if (h.empty()) h.assign(static_cast<size_t>(nv), 0);
  int64_t run = 0;
  for (int64_t v = 0; v < nv; ++v) {
    offsets[v] = run;
    for (int t = 0; t < threads; ++t) { const int64_t c = hist[t][v]; hist[t][v] = run; run += c; }
  }
  offsets[nv] = run;
  parallel_for(ne, threads, [&](int t, int64_t lo, int64_t hi) {
    (void)per;
    for (int64_t e = lo; e < hi; ++e) indices[hist[t][dst[e]]++] = src[e];
  });
}

}  // extern "C"
