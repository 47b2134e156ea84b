// TEST INFRASTRUCTURE ONLY. C entry points around the reference's own host references for the
// graph structure, compiled IN PLACE (build_ref.sh extracts the function templates into
// temporary .inc files; nothing of them is stored in this repository):
//   graph_reference        /root/reference/cpp/tests/structure/graph_test.cpp:38-69
//   degree_reference       /root/reference/cpp/tests/structure/degree_test.cpp:37-57
//   weight_sum_reference   /root/reference/cpp/tests/structure/weight_sum_test.cpp:37-59
// Used by tests/ to pin oracle/graph_oracle.py (compressed build, degrees, weight sums).
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <optional>
#include <tuple>
#include <vector>

#include "ref_graph.inc"
#include "ref_degree.inc"
#include "ref_weight_sum.inc"

extern "C" {

// COO -> compressed rows (by destination when store_transposed), neighbours in input order
long ref_graph_i32_f64(int store_transposed, const int32_t* src, const int32_t* dst,
                       const double* weights /* or NULL */, int32_t num_vertices, int64_t num_edges,
                       int64_t* offsets /* V + 1 */, int32_t* indices /* E */, double* out_weights)
{
  auto w = weights ? std::make_optional<double const*>(weights) : std::nullopt;
  auto r = store_transposed
             ? graph_reference<true, int32_t, int64_t, double>(src, dst, w, num_vertices, num_edges)
             : graph_reference<false, int32_t, int64_t, double>(src, dst, w, num_vertices, num_edges);
  std::copy(std::get<0>(r).begin(), std::get<0>(r).end(), offsets);
  std::copy(std::get<1>(r).begin(), std::get<1>(r).end(), indices);
  if (weights) std::copy(std::get<2>(r)->begin(), std::get<2>(r)->end(), out_weights);
  return 0;
}

long ref_degree_i32(const int64_t* offsets, const int32_t* indices, int64_t* degrees,
                    int32_t num_vertices, int major)
{
  degree_reference<int32_t, int64_t>(offsets, indices, degrees, num_vertices, major != 0);
  return 0;
}

long ref_weight_sum_i32_f64(const int64_t* offsets, const int32_t* indices, const double* weights,
                            double* sums, int32_t num_vertices, int major)
{
  weight_sum_reference<int32_t, int64_t, double>(offsets, indices, weights, sums, num_vertices,
                                                 major != 0);
  return 0;
}

}  // extern "C"
