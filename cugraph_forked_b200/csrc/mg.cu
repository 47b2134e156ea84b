// Multi-GPU (2D-partitioned) graph build and drivers.  Placeholder until the MG path lands.
#include "centrality.cuh"

namespace b200 {

void pagerank_mg(b200_handle*, b200_graph*, const b200_pagerank_params_t*, void*,
                 b200_centrality_stats_t*)
{
  throw error(kNotImplemented, "multi-GPU PageRank is not implemented yet");
}
void katz_mg(b200_handle*, b200_graph*, const b200_katz_params_t*, void*, b200_centrality_stats_t*)
{
  throw error(kNotImplemented, "multi-GPU Katz centrality is not implemented yet");
}
void graph_create_mg(b200_handle*, const void*, const void*, const void*, size_t, int, int,
                     const void*, size_t, bool, b200_graph*)
{
  throw error(kNotImplemented, "multi-GPU graph creation is not implemented yet");
}

}  // namespace b200
