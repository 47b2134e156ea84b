/*
 * cugraph_c/graph.h - drop-in for cpp/include/cugraph_c/graph.h (implementation in the
 * reference: cpp/src/c_api/graph_sg.cpp, graph_mg.cpp).  Edge ids / edge types are accepted
 * and ignored by this library: the pull path (PageRank, Katz, eigenvector) never reads them.
 */
#pragma once

#include <cugraph_c/array.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t align_;
} cugraph_graph_t;

typedef struct {
  bool_t is_symmetric;
  bool_t is_multigraph;
} cugraph_graph_properties_t;

/* Single-GPU graph from a device COO.  renumber = FALSE requires ids 0..V-1.
 * The edge type is INT64 iff the edge count is >= 2^31 - 1 (graph_sg.cpp:254-292). */
cugraph_error_code_t cugraph_sg_graph_create(
  const cugraph_resource_handle_t* handle,
  const cugraph_graph_properties_t* properties,
  const cugraph_type_erased_device_array_view_t* src,
  const cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* weights,
  const cugraph_type_erased_device_array_view_t* edge_ids,
  const cugraph_type_erased_device_array_view_t* edge_types,
  bool_t store_transposed,
  bool_t renumber,
  bool_t check,
  cugraph_graph_t** graph,
  cugraph_error_t** error);

void cugraph_sg_graph_free(cugraph_graph_t* graph);

/* Multi-GPU graph: every rank passes the edges it holds; they are shuffled to their 2D owner
 * and the graph is always renumbered (graph_mg.cpp:238-378). num_edges = global edge count. */
cugraph_error_code_t cugraph_mg_graph_create(
  const cugraph_resource_handle_t* handle,
  const cugraph_graph_properties_t* properties,
  const cugraph_type_erased_device_array_view_t* src,
  const cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* weights,
  const cugraph_type_erased_device_array_view_t* edge_ids,
  const cugraph_type_erased_device_array_view_t* edge_types,
  bool_t store_transposed,
  size_t num_edges,
  bool_t check,
  cugraph_graph_t** graph,
  cugraph_error_t** error);

void cugraph_mg_graph_free(cugraph_graph_t* graph);

#ifdef __cplusplus
}
#endif
