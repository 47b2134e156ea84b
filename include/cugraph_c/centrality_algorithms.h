/*
 * cugraph_c/centrality_algorithms.h - the pull-path subset of
 * cpp/include/cugraph_c/centrality_algorithms.h: PageRank, personalized PageRank, Katz and
 * eigenvector centrality plus their result object (reference implementations:
 * cpp/src/c_api/pagerank.cpp:244-385, katz.cpp:146-162, eigenvector_centrality.cpp,
 * centrality_result.cpp:21-43).  HITS and betweenness are outside this library's scope.
 *
 * Ownership: the callee allocates *result (vertex ids = external ids in internal order, values =
 * scores in the same order) and *error; free them with cugraph_centrality_result_free /
 * cugraph_error_free.  The getters return non-owning views.  cugraph_pagerank may rebuild the
 * graph in place when it was created with store_transposed = FALSE, as the reference does.
 */
#pragma once

#include <cugraph_c/array.h>
#include <cugraph_c/error.h>
#include <cugraph_c/graph.h>
#include <cugraph_c/resource_handle.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t align_;
} cugraph_centrality_result_t;

cugraph_type_erased_device_array_view_t* cugraph_centrality_result_get_vertices(
  cugraph_centrality_result_t* result);

cugraph_type_erased_device_array_view_t* cugraph_centrality_result_get_values(
  cugraph_centrality_result_t* result);

void cugraph_centrality_result_free(cugraph_centrality_result_t* result);

cugraph_error_code_t cugraph_pagerank(
  const cugraph_resource_handle_t* handle,
  cugraph_graph_t* graph,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_vertices,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_sums,
  const cugraph_type_erased_device_array_view_t* initial_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_guess_values,
  double alpha,
  double epsilon,
  size_t max_iterations,
  bool_t do_expensive_check,
  cugraph_centrality_result_t** result,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_personalized_pagerank(
  const cugraph_resource_handle_t* handle,
  cugraph_graph_t* graph,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_vertices,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_sums,
  const cugraph_type_erased_device_array_view_t* initial_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_guess_values,
  const cugraph_type_erased_device_array_view_t* personalization_vertices,
  const cugraph_type_erased_device_array_view_t* personalization_values,
  double alpha,
  double epsilon,
  size_t max_iterations,
  bool_t do_expensive_check,
  cugraph_centrality_result_t** result,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_eigenvector_centrality(const cugraph_resource_handle_t* handle,
                                                    cugraph_graph_t* graph,
                                                    double epsilon,
                                                    size_t max_iterations,
                                                    bool_t do_expensive_check,
                                                    cugraph_centrality_result_t** result,
                                                    cugraph_error_t** error);

cugraph_error_code_t cugraph_katz_centrality(const cugraph_resource_handle_t* handle,
                                             cugraph_graph_t* graph,
                                             const cugraph_type_erased_device_array_view_t* betas,
                                             double alpha,
                                             double beta,
                                             double epsilon,
                                             size_t max_iterations,
                                             bool_t do_expensive_check,
                                             cugraph_centrality_result_t** result,
                                             cugraph_error_t** error);

#ifdef __cplusplus
}
#endif
