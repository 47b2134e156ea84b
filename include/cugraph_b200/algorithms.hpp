/*
 * cugraph_b200/algorithms.hpp - header-only C++ shims that keep the reference's C++ entry points
 * for the pull path alive on top of the raw C ABI (b200_cugraph.h):
 *
 *   cugraph::pagerank                 cpp/include/cugraph/algorithms.hpp:1147-1159
 *   cugraph::katz_centrality          cpp/include/cugraph/algorithms.hpp:1272-1283
 *   cugraph::eigenvector_centrality   cpp/include/cugraph/algorithms.hpp:1187-1193
 *
 * Same template parameters, argument order, defaults and error behaviour (a failed call throws
 * cugraph::logic_error carrying the library's message).  raft and rmm are not available to this
 * library, so two minimal stand-ins carry what the signatures need:
 *   raft::handle_t      -> cugraph::b200::handle_t   (owns or borrows a b200_handle_t*)
 *   graph_view_t<...>   -> cugraph::graph_view_t     (non-owning view of a b200_graph_t*)
 * A libcugraph build that links this library instead of its own kernels keeps its
 * raft::handle_t / graph_t and forwards `handle.get_stream()` and the graph's device arrays
 * through b200_handle_create(stream) / b200_graph_create_sg(); see INTEGRATION.md.
 */
#pragma once

#include <b200_cugraph.h>

#include <cstddef>
#include <cstdint>
#include <optional>
#include <tuple>
#include <stdexcept>
#include <string>
#include <type_traits>

namespace cugraph {

struct logic_error : public std::logic_error {
  explicit logic_error(std::string const& m) : std::logic_error(m) {}
};

namespace b200 {

inline void check(int code, char const* what)
{
  if (code != 0) throw logic_error(std::string(what) + ": " + b200_last_error());
}

template <typename T>
constexpr int dtype_id()
{
  if constexpr (std::is_same_v<T, int32_t>) return B200_INT32;
  else if constexpr (std::is_same_v<T, int64_t>) return B200_INT64;
  else if constexpr (std::is_same_v<T, float>) return B200_FLOAT32;
  else {
    static_assert(std::is_same_v<T, double>, "unsupported type");
    return B200_FLOAT64;
  }
}

// stand-in for raft::handle_t: one stream (+ NCCL communicators when created with an id)
class handle_t {
 public:
  explicit handle_t(void* cuda_stream = nullptr) { check(b200_handle_create(cuda_stream, &h_), "handle"); }
  handle_t(void* cuda_stream, void const* nccl_unique_id, int rank, int world_size, int row_comm_size)
  {
    check(b200_handle_create_mg(cuda_stream, nccl_unique_id, rank, world_size, row_comm_size, &h_),
          "handle");
  }
  handle_t(handle_t const&)            = delete;
  handle_t& operator=(handle_t const&) = delete;
  ~handle_t() { b200_handle_free(h_); }
  b200_handle_t* get() const { return h_; }

 private:
  b200_handle_t* h_{nullptr};
};

}  // namespace b200

// non-owning view; the template parameters are the reference's (graph_view.hpp:258-262)
template <typename vertex_t, typename edge_t, typename weight_t, bool store_transposed, bool multi_gpu>
class graph_view_t {
 public:
  explicit graph_view_t(b200_graph_t* g) : g_(g) {}
  b200_graph_t* get() const { return g_; }
  vertex_t number_of_vertices() const { return static_cast<vertex_t>(view().num_vertices); }
  edge_t number_of_edges() const { return static_cast<edge_t>(view().num_edges); }
  vertex_t local_vertex_partition_range_first() const
  {
    return static_cast<vertex_t>(view().local_vertex_first);
  }
  vertex_t local_vertex_partition_range_size() const
  {
    auto v = view();
    return static_cast<vertex_t>(v.local_vertex_last - v.local_vertex_first);
  }

 private:
  b200_graph_view_t view() const
  {
    b200_graph_view_t v;
    b200::check(b200_graph_view(g_, &v), "graph view");
    return v;
  }
  b200_graph_t* g_;
};

template <typename vertex_t, typename edge_t, typename weight_t, typename result_t, bool multi_gpu>
void pagerank(b200::handle_t const& handle,
              graph_view_t<vertex_t, edge_t, weight_t, true, multi_gpu> const& graph_view,
              std::optional<weight_t const*> precomputed_vertex_out_weight_sums,
              std::optional<vertex_t const*> personalization_vertices,
              std::optional<result_t const*> personalization_values,
              std::optional<vertex_t> personalization_vector_size,
              result_t* pageranks,
              result_t alpha,
              result_t epsilon,
              size_t max_iterations   = 500,
              bool has_initial_guess  = false,
              bool do_expensive_check = false)
{
  (void)do_expensive_check;
  static_assert(std::is_same_v<vertex_t, int32_t>,
                "internal vertex ids are 32-bit (V < 2^31); external int64 ids are handled by "
                "the renumber map");
  if (precomputed_vertex_out_weight_sums && !std::is_same_v<weight_t, result_t>)
    throw logic_error("precomputed out-weight sums must have the result type");
  // the reference's own argument checks (pagerank_impl.cuh:78-93)
  if (personalization_vertices.has_value() &&
      !(personalization_values.has_value() && personalization_vector_size.has_value()))
    throw logic_error(
      "Invalid input argument: if personalization_vertices.has_value() is true, "
      "personalization_values.has_value() and personalization_vector_size.has_value() should be "
      "true as well.");
  b200_pagerank_params_t p{};
  p.alpha             = static_cast<double>(alpha);
  p.epsilon           = static_cast<double>(epsilon);
  p.max_iterations    = max_iterations;
  p.value_type        = b200::dtype_id<result_t>();
  p.has_initial_guess = has_initial_guess ? 1 : 0;
  p.precomputed_out_weight_sums =
    precomputed_vertex_out_weight_sums ? static_cast<void const*>(*precomputed_vertex_out_weight_sums)
                                       : nullptr;
  if (personalization_vertices) {
    p.personalization_vertices = *personalization_vertices;
    p.personalization_values   = *personalization_values;
    p.personalization_size     = static_cast<uint64_t>(*personalization_vector_size);
  }
  b200::check(b200_pagerank(handle.get(), graph_view.get(), &p, pageranks, nullptr), "pagerank");
}

template <typename vertex_t, typename edge_t, typename weight_t, typename result_t, bool multi_gpu>
void katz_centrality(b200::handle_t const& handle,
                     graph_view_t<vertex_t, edge_t, weight_t, true, multi_gpu> const& graph_view,
                     result_t const* betas,
                     result_t* katz_centralities,
                     result_t alpha,
                     result_t beta,
                     result_t epsilon,
                     size_t max_iterations   = 500,
                     bool has_initial_guess  = false,
                     bool normalize          = false,
                     bool do_expensive_check = false)
{
  (void)do_expensive_check;
  b200_katz_params_t p{};
  p.alpha             = static_cast<double>(alpha);
  p.beta              = static_cast<double>(beta);
  p.epsilon           = static_cast<double>(epsilon);
  p.max_iterations    = max_iterations;
  p.value_type        = b200::dtype_id<result_t>();
  p.has_initial_guess = has_initial_guess ? 1 : 0;
  p.normalize         = normalize ? 1 : 0;
  p.betas             = betas;
  b200::check(b200_katz_centrality(handle.get(), graph_view.get(), &p, katz_centralities, nullptr),
              "katz_centrality");
}

// The reference returns an rmm::device_uvector<weight_t>; without rmm the caller passes the
// output array (local vertex partition size).  `initial_centralities` == nullptr starts from
// 1 / V as the reference does.
template <typename vertex_t, typename edge_t, typename weight_t, bool multi_gpu>
void eigenvector_centrality(b200::handle_t const& handle,
                            graph_view_t<vertex_t, edge_t, weight_t, true, multi_gpu> const& graph_view,
                            bool has_initial_centralities,
                            weight_t* centralities,
                            weight_t epsilon,
                            size_t max_iterations   = 500,
                            bool do_expensive_check = false)
{
  (void)do_expensive_check;
  b200_eigenvector_params_t p{};
  p.epsilon           = static_cast<double>(epsilon);
  p.max_iterations    = max_iterations;
  p.value_type        = b200::dtype_id<weight_t>();
  p.has_initial_guess = has_initial_centralities ? 1 : 0;
  b200::check(b200_eigenvector_centrality(handle.get(), graph_view.get(), &p, centralities, nullptr),
              "eigenvector_centrality");
}

// cugraph::hits (cpp/include/cugraph/algorithms.hpp:1224-1234): returns (sum of the differences of
// the hub scores of the last two iterations, iteration count).  Single-GPU graphs.
template <typename vertex_t, typename edge_t, typename weight_t, bool multi_gpu>
std::tuple<weight_t, size_t> hits(b200::handle_t const& handle,
                                  graph_view_t<vertex_t, edge_t, weight_t, true, multi_gpu> const& graph_view,
                                  weight_t* hubs,
                                  weight_t* authorities,
                                  weight_t epsilon,
                                  size_t max_iterations,
                                  bool has_initial_hubs_guess,
                                  bool normalize,
                                  bool do_expensive_check)
{
  (void)do_expensive_check;
  b200_hits_params_t p{};
  p.epsilon                = static_cast<double>(epsilon);
  p.max_iterations         = max_iterations;
  p.value_type             = b200::dtype_id<weight_t>();
  p.has_initial_hubs_guess = has_initial_hubs_guess ? 1 : 0;
  p.normalize              = normalize ? 1 : 0;
  b200_centrality_stats_t st{};
  b200::check(b200_hits(handle.get(), graph_view.get(), &p, hubs, authorities, &st), "hits");
  return std::make_tuple(static_cast<weight_t>(st.last_diff), static_cast<size_t>(st.iterations));
}

}  // namespace cugraph
