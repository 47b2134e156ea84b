// TEST INFRASTRUCTURE ONLY — C-ABI shim around the reference's OWN host oracles.
//
// oracle/build_ref.sh pulls the three self-contained function templates out of the
// reference's gtest sources (where they lie under /root/reference; nothing is copied
// into this repository) into a temporary include directory and compiles this shim
// against them, producing oracle/_ref/libcugraph_ref_oracle.so (git-ignored).
//   ref_pagerank.inc     <- cpp/tests/link_analysis/pagerank_test.cpp:42-130
//   ref_katz.inc         <- cpp/tests/centrality/katz_centrality_test.cpp:41-93
//   ref_eigenvector.inc  <- cpp/tests/centrality/eigenvector_centrality_test.cpp:41-96
//   ref_hits.inc         <- cpp/tests/link_analysis/hits_test.cpp:46-127
// The only adaptation is the gtest macro: ASSERT_TRUE(x) throws, and the shim turns
// the exception into a -1 return (the reference aborts the test instead).
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <numeric>
#include <optional>
#include <stdexcept>
#include <tuple>
#include <vector>

#define ASSERT_TRUE(cond)                                           \
  do {                                                              \
    if (!(cond)) throw std::runtime_error("ASSERT_TRUE(" #cond ")"); \
  } while (0)

#define CUGRAPH_EXPECTS(cond, msg)                 \
  do {                                             \
    if (!(cond)) throw std::runtime_error(msg);    \
  } while (0)

#include "ref_pagerank.inc"
#include "ref_katz.inc"
#include "ref_eigenvector.inc"
#include "ref_hits.inc"

extern "C" {

#define REF_DEFINE(SUFFIX, R, E)                                                                  \
  long ref_pagerank_##SUFFIX(const E* offsets, const int32_t* indices, const R* weights,          \
                             const int32_t* pers_vertices, const R* pers_values,                  \
                             int64_t pers_size, R* pr, int64_t nv, double alpha, double epsilon,  \
                             long max_iterations, int has_initial_guess)                          \
  {                                                                                               \
    try {                                                                                         \
      pagerank_reference<int32_t, E, R, R>(                                                       \
        offsets, indices, weights ? std::optional<const R*>{weights} : std::nullopt,              \
        pers_vertices ? std::optional<const int32_t*>{pers_vertices} : std::nullopt,              \
        pers_vertices ? std::optional<const R*>{pers_values} : std::nullopt,                      \
        pers_vertices ? std::optional<int32_t>{static_cast<int32_t>(pers_size)} : std::nullopt,   \
        pr, static_cast<int32_t>(nv), static_cast<R>(alpha), static_cast<R>(epsilon),             \
        static_cast<size_t>(max_iterations), has_initial_guess != 0);                             \
    } catch (std::exception const&) {                                                             \
      return -1;                                                                                  \
    }                                                                                             \
    return 0;                                                                                     \
  }                                                                                               \
  long ref_katz_##SUFFIX(const E* offsets, const int32_t* indices, const R* weights,              \
                         const R* betas, R* kc, int64_t nv, double alpha, double beta,            \
                         double epsilon, long max_iterations, int has_initial_guess,              \
                         int normalize)                                                           \
  {                                                                                               \
    try {                                                                                         \
      katz_centrality_reference<int32_t, E, R, R>(                                                \
        offsets, indices, weights ? std::optional<const R*>{weights} : std::nullopt, betas, kc,   \
        static_cast<int32_t>(nv), static_cast<R>(alpha), static_cast<R>(beta),                    \
        static_cast<R>(epsilon), static_cast<size_t>(max_iterations), has_initial_guess != 0,     \
        normalize != 0);                                                                          \
    } catch (std::exception const&) {                                                             \
      return -1;                                                                                  \
    }                                                                                             \
    return 0;                                                                                     \
  }

REF_DEFINE(f32, float, int64_t)
REF_DEFINE(f64, double, int64_t)

#define REF_DEFINE_EIG(SUFFIX, R)                                                                \
  long ref_eigenvector_##SUFFIX(const int32_t* src, const int32_t* dst, const R* weights,        \
                                int64_t ne, R* out, int64_t nv, double epsilon,                  \
                                long max_iterations)                                             \
  {                                                                                              \
    try {                                                                                        \
      eigenvector_centrality_reference<int32_t, R>(                                              \
        src, dst, weights ? std::optional<const R*>{weights} : std::nullopt,                     \
        static_cast<size_t>(ne), out, static_cast<int32_t>(nv), static_cast<R>(epsilon),         \
        static_cast<size_t>(max_iterations));                                                    \
    } catch (std::exception const&) {                                                            \
      return -1;                                                                                 \
    }                                                                                            \
    return 0;                                                                                    \
  }

REF_DEFINE_EIG(f32, float)
REF_DEFINE_EIG(f64, double)

// hits_reference always applies the final sum normalisation; returns its iteration count
#define REF_DEFINE_HITS(SUFFIX, R)                                                               \
  long ref_hits_##SUFFIX(const int64_t* offsets, const int32_t* indices, int64_t nv, int64_t ne,  \
                         long max_iterations, const R* starting_hubs, double tolerance, R* hubs,  \
                         R* authorities, double* error)                                           \
  {                                                                                              \
    try {                                                                                        \
      auto r = hits_reference<R, int32_t, int64_t>(                                              \
        offsets, indices, static_cast<int32_t>(nv), ne, static_cast<size_t>(max_iterations),     \
        starting_hubs ? std::optional<const R*>{starting_hubs} : std::nullopt, true, tolerance); \
      std::copy(std::get<0>(r).begin(), std::get<0>(r).end(), hubs);                             \
      std::copy(std::get<1>(r).begin(), std::get<1>(r).end(), authorities);                      \
      if (error) *error = std::get<2>(r);                                                        \
      return static_cast<long>(std::get<3>(r));                                                  \
    } catch (std::exception const&) {                                                            \
      return -1;                                                                                 \
    }                                                                                            \
  }

REF_DEFINE_HITS(f32, float)
REF_DEFINE_HITS(f64, double)

}  // extern "C"
