// TEST INFRASTRUCTURE ONLY — CPU oracle for the PageRank / Katz / eigenvector / HITS hot path.
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
//   hits_reference              /root/reference/cpp/tests/link_analysis/hits_test.cpp:46-127
//
// Differences from the reference (none change a converged result):
//   * non-convergence returns -1 instead of tripping a gtest ASSERT;
//   * `fixed_iterations > 0` runs exactly that many iterations with no stop test
//     (needed to compare fp64 scores at an identical iteration count);
//   * the executed iteration count is returned.
// Pinned by tests/test_oracle.py against the C-API golden vectors
// (cpp/tests/c_api/pagerank_test.c:211-318, katz_test.c:103-129,
// eigenvector_centrality_test.c:93-108, hits_test.c) and against oracle/_ref (the reference's own
// functions compiled from /root/reference by oracle/build_ref.sh).
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <vector>

namespace {

// A = accumulator type of the row sums and of the dangling / difference sums.  A = R restates
// pagerank_reference exactly (bit-identical, tests/test_oracle.py).  A = long double (x87 80-bit,
// 64-bit mantissa) is the "extended" variant: the same algorithm with the SUMS kept in extended
// precision and every stored value still an R.  The sequential sums of the reference carry up to
// deg * eps of rounding error on hub rows (in-degree > 1e5 on R-MAT scale 20 => ~4e-12 in fp64,
// ~3e-3 in fp32); the extended variant is what both the reference and a device implementation
// with another summation order are measured against in tests/test_gpu_configs.py.
template <typename R, typename A = R>
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
    A dangling_acc{0.0};
    for (int64_t v = 0; v < nv; ++v)
      if (outw[v] == R{0.0}) dangling_acc += old[v];
    const R dangling = static_cast<R>(dangling_acc);
    for (int64_t v = 0; v < nv; ++v) {
      A acc{0.0};
      for (int64_t e = offsets[v]; e < offsets[v + 1]; ++e) {
        int32_t u = indices[e];
        R w = weights ? weights[e] : R{1.0};
        acc += alpha * old[u] * (w / outw[u]);
      }
      pr[v] = static_cast<R>(acc);
      if (!pers_vertices) pr[v] += (dangling * alpha + (1.0 - alpha)) / static_cast<R>(nv);
    }
    if (pers_vertices) {
      for (int64_t i = 0; i < pers_size; ++i)
        pr[pers_vertices[i]] += (dangling * alpha + (1.0 - alpha)) * (pers_values[i] / pers_sum);
    }
    A diff_acc{0.0};
    for (int64_t v = 0; v < nv; ++v) diff_acc += std::abs(pr[v] - old[v]);
    const R diff = static_cast<R>(diff_acc);
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

// ------------------------------------------------------------------------------------------
// HITS (the next row of the path).  Restates the reference's host
// hits_reference (cpp/tests/link_analysis/hits_test.cpp:46-127) with the two switches of the
// device implementation it checks (cpp/src/link_analysis/hits_impl.cuh:54-185): the final sum
// normalisation is optional (`normalize`, hits_impl.cuh:166-169; the host function always
// applies it) and the iteration count can be pinned (`fixed_iterations`).  offsets / indices
// are the CSR of the graph (rows = sources); weights play no role in HITS.
// Returns the 0-based index of the iteration that met the tolerance, or max_iterations.
// ------------------------------------------------------------------------------------------
// The reference sums with std::reduce / std::transform_reduce, which libstdc++ evaluates for
// random-access ranges four elements at a time: init += (a0 + a1) + (a2 + a3).  Restated here so
// that the port stays bit-identical to the compiled reference (tests/test_oracle.py).
template <typename R, typename F>
R reduce4(int64_t n, F get)
{
  R init = R(0);
  int64_t i = 0;
  while (n - i >= 4) {
    const R v1 = get(i) + get(i + 1);
    const R v2 = get(i + 2) + get(i + 3);
    const R v3 = v1 + v2;
    init       = init + v3;
    i += 4;
  }
  for (; i < n; ++i) init = init + get(i);
  return init;
}

template <typename R>
long hits_cpu(const int64_t* offsets, const int32_t* indices, int64_t nv, long max_iterations,
              const R* starting_hubs, int normalize, double tolerance, R* hubs_out, R* auth_out,
              double* error_out, long fixed_iterations)
{
  if (nv < 2) return -2;
  std::vector<R> prev_hubs(nv, R(1) / static_cast<R>(nv)), curr_hubs(nv), curr_auth(nv);
  if (starting_hubs) {
    std::copy(starting_hubs, starting_hubs + nv, prev_hubs.begin());
    const R norm = reduce4<R>(nv, [&](int64_t v) { return prev_hubs[v]; });
    for (int64_t v = 0; v < nv; ++v) prev_hubs[v] = prev_hubs[v] / norm;
  }
  double err = std::numeric_limits<double>::max();
  const long cap = fixed_iterations > 0 ? fixed_iterations : max_iterations;
  long it = 0;
  for (; it < cap; ++it) {
    std::fill(curr_hubs.begin(), curr_hubs.end(), R(0));
    std::fill(curr_auth.begin(), curr_auth.end(), R(0));
    for (int64_t s = 0; s < nv; ++s)
      for (int64_t e = offsets[s]; e < offsets[s + 1]; ++e) curr_auth[indices[e]] += prev_hubs[s];
    for (int64_t s = 0; s < nv; ++s)
      for (int64_t e = offsets[s]; e < offsets[s + 1]; ++e) curr_hubs[s] += curr_auth[indices[e]];
    R hmax = R(0), amax = R(0);
    for (int64_t v = 0; v < nv; ++v) hmax = std::max(hmax, curr_hubs[v]);
    for (int64_t v = 0; v < nv; ++v) curr_hubs[v] = curr_hubs[v] / hmax;
    for (int64_t v = 0; v < nv; ++v) amax = std::max(amax, curr_auth[v]);
    for (int64_t v = 0; v < nv; ++v) curr_auth[v] = curr_auth[v] / amax;
    const R diff =
      reduce4<R>(nv, [&](int64_t v) { return std::abs(curr_hubs[v] - prev_hubs[v]); });
    err = static_cast<double>(diff);
    if (fixed_iterations <= 0 && err < tolerance) break;
    std::copy(curr_hubs.begin(), curr_hubs.end(), prev_hubs.begin());
  }
  if (normalize) {
    const R hs = reduce4<R>(nv, [&](int64_t v) { return curr_hubs[v]; });
    for (int64_t v = 0; v < nv; ++v) curr_hubs[v] = curr_hubs[v] / hs;
    const R as = reduce4<R>(nv, [&](int64_t v) { return curr_auth[v]; });
    for (int64_t v = 0; v < nv; ++v) curr_auth[v] = curr_auth[v] / as;
  }
  std::copy(curr_hubs.begin(), curr_hubs.end(), hubs_out);
  std::copy(curr_auth.begin(), curr_auth.end(), auth_out);
  if (error_out) *error_out = err;
  return it;
}

}  // namespace

extern "C" {

#define ORACLE_DEFINE_HITS(SUFFIX, R)                                                              \
  long oracle_hits_##SUFFIX(const int64_t* offsets, const int32_t* indices, int64_t nv,            \
                            long max_iterations, const R* starting_hubs, int normalize,            \
                            double tolerance, R* hubs, R* authorities, double* error,              \
                            long fixed_iterations)                                                 \
  {                                                                                                \
    return hits_cpu<R>(offsets, indices, nv, max_iterations, starting_hubs, normalize, tolerance,  \
                       hubs, authorities, error, fixed_iterations);                                \
  }
ORACLE_DEFINE_HITS(f32, float)
ORACLE_DEFINE_HITS(f64, double)


extern "C" long oracle_pagerank_f64_ext(const int64_t* offsets, const int32_t* indices,
                                        const double* weights, const int32_t* pers_vertices,
                                        const double* pers_values, int64_t pers_size, double* pr,
                                        int64_t nv, double alpha, double epsilon,
                                        long max_iterations, int has_initial_guess,
                                        long fixed_iterations)
{
  return pagerank_cpu<double, long double>(offsets, indices, weights, pers_vertices, pers_values,
                                           pers_size, pr, nv, alpha, epsilon, max_iterations,
                                           has_initial_guess, fixed_iterations);
}

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
