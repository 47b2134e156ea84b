/*
 * cugraph_c/centrality_algorithms.h - the pull-path subset of
 * cpp/include/cugraph_c/centrality_algorithms.h: PageRank, personalized PageRank, Katz and
 * eigenvector centrality plus their result object (reference implementations:
 * cpp/src/c_api/pagerank.cpp:244-385, katz.cpp:146-162, eigenvector_centrality.cpp,
 * centrality_result.cpp:21-43), and HITS with its own result object (cpp/src/c_api/hits.cpp:34-245).
 * Betweenness is outside this library's scope.
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

/* ---- HITS (cpp/include/cugraph_c/centrality_algorithms.h:235-332; c_api/hits.cpp) -------------
 * Single-GPU graphs.  The graph may be rebuilt in place when it was created with
 * store_transposed = FALSE (hits.cpp:94-100).  initial_hubs_guess_* are (external vertex id,
 * value) pairs or NULL; vertices without a pair start at 0 (hits.cpp:113-139). */
typedef struct {
  int32_t align_;
} cugraph_hits_result_t;

cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_vertices(
  cugraph_hits_result_t* result);
cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_hubs(
  cugraph_hits_result_t* result);
cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_authorities(
  cugraph_hits_result_t* result);
double cugraph_hits_result_get_hub_score_differences(cugraph_hits_result_t* result);
size_t cugraph_hits_result_get_number_of_iterations(cugraph_hits_result_t* result);
void cugraph_hits_result_free(cugraph_hits_result_t* result);

cugraph_error_code_t cugraph_hits(
  const cugraph_resource_handle_t* handle,
  cugraph_graph_t* graph,
  double epsilon,
  size_t max_iterations,
  const cugraph_type_erased_device_array_view_t* initial_hubs_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_hubs_guess_values,
  bool_t normalize,
  bool_t do_expensive_check,
  cugraph_hits_result_t** result,
  cugraph_error_t** error);

#ifdef __cplusplus
}
#endif
