// Internal declarations shared by graph.cu, centrality.cu, mg.cu and api.cu.
#pragma once

#include <algorithm>

#include "common.cuh"
#include "prims.cuh"

namespace b200 {

// cpp/include/cugraph/graph_view.hpp:255-263
constexpr int64_t kLowDegreeThreshold  = 32;
constexpr int64_t kMidDegreeThreshold  = 1024;
constexpr double kHypersparseRatio     = 0.5;

void graph_create_sg(b200_handle* h, const void* src, const void* dst, const void* weights,
                     size_t ne, int vertex_type, int weight_type, const void* vertices,
                     size_t n_given, bool store_transposed, bool renumber, b200_graph* g);
void graph_out_weight_sums(b200_handle* h, b200_graph* g);
void graph_in_degrees(b200_handle* h, const b200_graph* g, void* out);
// out[i] = internal (local) id of external id ext_ids[i], or -1
void graph_ext_to_internal(b200_handle* h, b200_graph* g, const void* ext_ids, size_t n,
                           int32_t* out);
void graph_transpose_storage(b200_handle* h, b200_graph* g);
// multi-GPU counterpart (graph_mg.cu): local histograms + one reduction per row_comm peer
void graph_out_weight_sums_mg(b200_handle* h, b200_graph* g);
// helpers shared by the single- and multi-GPU builds (graph.cu)
void launch_widen_ids(const void* a, const void* b, size_t n, int width, uint64_t* out,
                      cudaStream_t s);
int64_t launch_unique_ids(const uint64_t* sorted, size_t n, int width, void* out, cudaStream_t s);
// (re)build the tile schedule of one edge partition: tiles of <= T edges and <= rowcap rows
void build_schedule(b200_handle* h, edge_partition& ep, int T, int rowcap);

}  // namespace b200
