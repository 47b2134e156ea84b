// Internal declarations of the drivers (centrality.cu single GPU, mg.cu multi-GPU).
#pragma once

#include "epilogues.cuh"
#include "graph.cuh"

namespace b200 {

void pagerank(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, void* pageranks,
              b200_centrality_stats_t* stats);
void katz_centrality(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, void* out,
                     b200_centrality_stats_t* stats);
void eigenvector_centrality(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p,
                            void* out, b200_centrality_stats_t* stats);
void spmv_pull(b200_handle* h, b200_graph* g, int value_type, const void* x, double scale,
               double init, void* y);
// hits.cu
void hits(b200_handle* h, b200_graph* g, const b200_hits_params_t* p, void* hubs, void* authorities,
          b200_centrality_stats_t* stats);

// one edge partition's SpMV with the raw epilogue (centrality.cu), used by the multi-GPU drivers
struct partition_spmv;
partition_spmv* partition_spmv_create(b200_handle* h, b200_graph* g, int part, int value_type);
void partition_spmv_run(partition_spmv* p, const void* x, void* y, const iter_state* state);
// the next partition_spmv_run may start with the minors [lo, hi) of x; the rest is valid once
// `rest_ready` has fired (ignored by layouts that cannot split their first phase)
void partition_spmv_split(partition_spmv* p, int64_t lo, int64_t hi, cudaEvent_t rest_ready);
int64_t partition_spmv_launches(const partition_spmv* p);
void partition_spmv_free(partition_spmv* p);

// mg.cu
void eigenvector_mg(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p, void* out,
                    b200_centrality_stats_t* stats);
void pagerank_mg(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, void* pageranks,
                 b200_centrality_stats_t* stats);
void katz_mg(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, void* out,
             b200_centrality_stats_t* stats);
// graph_mg.cu
void graph_create_mg(b200_handle* h, const void* src, const void* dst, const void* weights,
                     size_t ne, int vertex_type, int weight_type, const void* local_vertices,
                     size_t n_local_vertices, bool store_transposed, b200_graph* g);

// rmat.cu
void generate_rmat(b200_handle* h, int scale, size_t num_edges, double a, double b, double c,
                   uint64_t seed, bool clip_and_flip, uint64_t first_edge, int vertex_type,
                   void* src, void* dst);
void uniform_fill(b200_handle* h, int value_type, void* out, size_t n, uint64_t seed);

}  // namespace b200
