// TEST INFRASTRUCTURE ONLY — CPU oracle for the PageRank / Katz / eigenvector hot path.
//
// Plain C++17 restatement of the reference's host-side oracles. Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library; the product path (cugraph_forked_b200/csrc) never links it.
//
// Follows (arithmetic order kept, so results are bit-identical to the reference
// host oracle for the same CSC input):
//   pagerank_reference          /root/reference/cpp/tests/link_analysis/pagerank_test.cpp:42-130
//   katz_centrality_reference   /root/reference/cpp/tests/centrality/katz_centrality_test.cpp:41-93
//   eigenvector_..._reference   /root/reference/cpp/tests/centrality/eigenvector_centrality_test.cpp:41-96
//
// Differences from the reference (none change a converged result):
//   * non-convergence returns -1 instead of tripping a gtest ASSERT;
//   * `fixed_iterations > 0` runs exactly that many iterations with no stop test
//     (needed to compare fp64 scores at an identical iteration count);
//   * the executed iteration count is returned.
// Pinned by tests/test_oracle.py against the C-API golden vectors
// (cpp/tests/c_api/pagerank_test.c:211-318, katz_test.c:103-129,
// eigenvector_centrality_test.c:93-108) and against oracle/_ref (the reference's own
// functions compiled from /root/reference by oracle/build_ref.sh).
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace {

template <typename R>
long pagerank_cpu(const int64_t* offsets, const int32_t* indices, const R* weights,
                  const int32_t* pers_vertices, const R* pers_values, int64_t pers_size,
                  R* pr, int64_t nv, R alpha, R epsilon, long max_iterations,
                  int has_initial_guess, long fixed_iterations)
{
  if (nv == 0) return 0;
  if (has_initial_guess) {
    double s = 0.0;
    for (int64_t v = 0; v < nv; ++v) s += pr[v];
    R sum = static_cast<R>(s);
    if (!(sum > 0.0)) return -2;
    for (int64_t v = 0; v < nv; ++v) pr[v] /= sum;
  } else {
    for (int64_t v = 0; v < nv; ++v) pr[v] = R{1.0} / static_cast<R>(nv);
  }
  R pers_sum{0.0};
  if (pers_vertices) {
    double s = 0.0;
    for (int64_t i = 0; i < pers_size; ++i) s += pers_values[i];
    pers_sum = static_cast<R>(s);
    if (!(pers_sum > 0.0)) return -2;
  }
  // out-weight sum of every source, accumulated in CSC traversal order
  std::vector<R> outw(nv, R{0.0});
  for (int64_t v = 0; v < nv; ++v)
    for (int64_t e = offsets[v]; e < offsets[v + 1]; ++e)
      outw[indices[e]] += weights ? weights[e] : R{1.0};

  std::vector<R> old(nv, R{0.0});
  long iter = 0;
  for (;;) {
    for (int64_t v = 0; v < nv; ++v) old[v] = pr[v];
    R dangling{0.0};
    for (int64_t v = 0; v < nv; ++v)
      if (outw[v] == R{0.0}) dangling += old[v];
    for (int64_t v = 0; v < nv; ++v) {
      pr[v] = R{0.0};
      for (int64_t e = offsets[v]; e < offsets[v + 1]; ++e) {
        int32_t u = indices[e];
        R w = weights ? weights[e] : R{1.0};
        pr[v] += alpha * old[u] * (w / outw[u]);
      }
      if (!pers_vertices) pr[v] += (dangling * alpha + (1.0 - alpha)) / static_cast<R>(nv);
    }
    if (pers_vertices) {
      for (int64_t i = 0; i < pers_size; ++i)
        pr[pers_vertices[i]] += (dangling * alpha + (1.0 - alpha)) * (pers_values[i] / pers_sum);
    }
    R diff{0.0};
    for (int64_t v = 0; v < nv; ++v) diff += std::abs(pr[v] - old[v]);
    if (fixed_iterations > 0) {
      if (++iter >= fixed_iterations) return iter;
      continue;
    }
    if (diff < epsilon) return iter + 1;
    ++iter;
    if (!(iter < max_iterations)) return -1;
  }
}

template <typename R>
long katz_cpu(const int64_t* offsets, const int32_t* indices, const R* weights, const R* betas,
              R* kc, int64_t nv, R alpha, R beta, R epsilon, long max_iterations,
              int has_initial_guess, int normalize, long fixed_iterations)
{
  if (nv == 0) return 0;
  if (!has_initial_guess)
    for (int64_t v = 0; v < nv; ++v) kc[v] = R{0.0};
  std::vector<R> old(nv, R{0.0});
  long iter = 0, done = 0;
  for (;;) {
    for (int64_t v = 0; v < nv; ++v) old[v] = kc[v];
    for (int64_t v = 0; v < nv; ++v) {
      kc[v] = betas ? betas[v] : beta;
      for (int64_t e = offsets[v]; e < offsets[v + 1]; ++e) {
        R w = weights ? weights[e] : R{1.0};
        kc[v] += alpha * old[indices[e]] * w;
      }
    }
    R diff{0.0};
    for (int64_t v = 0; v < nv; ++v) diff += std::abs(kc[v] - old[v]);
    if (fixed_iterations > 0) {
      if (++iter >= fixed_iterations) { done = iter; break; }
      continue;
    }
    if (diff < epsilon) { done = iter + 1; break; }
    ++iter;
    if (!(iter < max_iterations)) return -1;
  }
  if (normalize) {
    R ss{0.0};
    for (int64_t v = 0; v < nv; ++v) ss += kc[v] * kc[v];
    R l2 = std::sqrt(ss);
    for (int64_t v = 0; v < nv; ++v) kc[v] = kc[v] / l2;
  }
  return done;
}

// NB: the reference eigenvector oracle works on the COO list with src/dst swapped
// (tmp[src] += old[dst] * w) and keeps its state in double for every weight_t.
template <typename R>
long eigenvector_cpu(const int32_t* src, const int32_t* dst, const R* weights, int64_t ne,
                     R* out, int64_t nv, R epsilon, long max_iterations, long fixed_iterations)
{
  if (nv == 0) return 0;
  std::vector<double> cur(nv, 1.0 / static_cast<double>(nv)), old(nv, 0.0);
  long iter = 0, done = 0;
  for (;;) {
    old = cur;
    for (int64_t v = 0; v < nv; ++v) cur[v] = 0.0;
    for (int64_t e = 0; e < ne; ++e) {
      R w = weights ? weights[e] : R{1.0};
      cur[src[e]] += old[dst[e]] * w;
    }
    double ss = 0.0;
    for (int64_t v = 0; v < nv; ++v) ss += cur[v] * cur[v];
    double l2 = std::sqrt(ss);
    for (int64_t v = 0; v < nv; ++v) cur[v] = cur[v] / l2;
    double diff = 0.0;
    for (int64_t v = 0; v < nv; ++v) diff += std::abs(cur[v] - old[v]);
    if (fixed_iterations > 0) {
      if (++iter >= fixed_iterations) { done = iter; break; }
      continue;
    }
    if (diff < (nv * epsilon)) { done = iter + 1; break; }
    ++iter;
    if (!(iter < max_iterations)) return -1;
  }
  for (int64_t v = 0; v < nv; ++v) out[v] = static_cast<R>(cur[v]);
  return done;
}

}  // namespace

extern "C" {

#define ORACLE_DEFINE(SUFFIX, R)                                                                   \
  long oracle_pagerank_##SUFFIX(const int64_t* offsets, const int32_t* indices, const R* weights,  \
                                const int32_t* pers_vertices, const R* pers_values,                \
                                int64_t pers_size, R* pr, int64_t nv, double alpha,                \
                                double epsilon, long max_iterations, int has_initial_guess,        \
                                long fixed_iterations)                                             \
  {                                                                                                \
    return pagerank_cpu<R>(offsets, indices, weights, pers_vertices, pers_values, pers_size, pr,   \
                           nv, static_cast<R>(alpha), static_cast<R>(epsilon), max_iterations,     \
                           has_initial_guess, fixed_iterations);                                   \
  }                                                                                                \
  long oracle_katz_##SUFFIX(const int64_t* offsets, const int32_t* indices, const R* weights,      \
                            const R* betas, R* kc, int64_t nv, double alpha, double beta,          \
                            double epsilon, long max_iterations, int has_initial_guess,            \
                            int normalize, long fixed_iterations)                                  \
  {                                                                                                \
    return katz_cpu<R>(offsets, indices, weights, betas, kc, nv, static_cast<R>(alpha),            \
                       static_cast<R>(beta), static_cast<R>(epsilon), max_iterations,              \
                       has_initial_guess, normalize, fixed_iterations);                            \
  }                                                                                                \
  long oracle_eigenvector_##SUFFIX(const int32_t* src, const int32_t* dst, const R* weights,       \
                                   int64_t ne, R* out, int64_t nv, double epsilon,                 \
                                   long max_iterations, long fixed_iterations)                     \
  {                                                                                                \
    return eigenvector_cpu<R>(src, dst, weights, ne, out, nv, static_cast<R>(epsilon),             \
                              max_iterations, fixed_iterations);                                   \
  }

ORACLE_DEFINE(f32, float)
ORACLE_DEFINE(f64, double)

}  // extern "C"
