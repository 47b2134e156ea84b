/*
 * b200_cugraph.h — raw-pointer C ABI of the B200-native pull-SpMV / PageRank engine.
 *
 * This is the boundary a libcugraph maintainer binds: plain pointers and sizes, no C++ or
 * torch types.  Every entry point names the reference interface it stands in for (paths are
 * relative to the reference tree).  The cugraph_c-compatible layer (the headers under include/cugraph_c,
 * cugraph_forked_b200/csrc/c_api.cu) and the Python mirror of pylibcugraph both sit on top
 * of exactly these calls.
 *
 * Conventions
 *   - all array arguments are DEVICE pointers unless the name starts with h_;
 *   - every function returns 0 on success or a cugraph_error_code_t value
 *     (cpp/include/cugraph_c/error.h:25-33); b200_last_error() returns the message of the
 *     last failure on the calling thread (the reference throws cugraph::logic_error with the
 *     same texts, cpp/include/cugraph/utilities/error.hpp:33-49);
 *   - dtype ids are the reference's data_type_id_t (cpp/include/cugraph_c/resource_handle.h):
 *     INT32=0, INT64=1, FLOAT32=2, FLOAT64=3;
 *   - calls are synchronous from the caller's view (they return after the work on the
 *     handle's stream has completed), as the reference's are.
 */
#ifndef B200_CUGRAPH_H
#define B200_CUGRAPH_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_INT32 0
#define B200_INT64 1
#define B200_FLOAT32 2
#define B200_FLOAT64 3

/* -------------------------------------------------------------------------------------------
 * handle  — stands in for raft::handle_t (stream + comms), cpp/src/c_api/resource_handle.cpp:21-43
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_handle b200_handle_t;

/* stream = cudaStream_t (NULL: a private non-blocking stream is created). */
int b200_handle_create(void* stream, b200_handle_t** out);
/* Multi-GPU handle: one process per GPU.  nccl_unique_id = the 128-byte ncclUniqueId produced
 * by b200_nccl_unique_id() on rank 0 and distributed by the launcher (torch.distributed / MPI).
 * row_comm_size = R of the reference's R x C grid (partition_manager.hpp:70-113); the row and
 * column sub-communicators are split exactly as the reference does: row_comm colour =
 * rank / R, col_comm colour = rank % R. */
int b200_handle_create_mg(void* stream, const void* nccl_unique_id, int rank, int world_size,
                          int row_comm_size, b200_handle_t** out);
int b200_nccl_unique_id(void* out_128_bytes);
int b200_handle_rank(const b200_handle_t* h);       /* cugraph_resource_handle_get_rank */
int b200_handle_world_size(const b200_handle_t* h);
void b200_handle_free(b200_handle_t* h);
const char* b200_last_error(void);

/* -------------------------------------------------------------------------------------------
 * graph — stands in for graph_t<vertex_t,edge_t,weight_t,store_transposed,false> +
 * create_graph_from_edgelist (cpp/src/structure/create_graph_from_edgelist_impl.cuh:570-774):
 * renumber by descending major degree (renumber_edgelist_impl.cuh:96-452), COO -> CSC/CSR with
 * ascending neighbours (structure/detail/structure_utils.cuh:173-490), degree segments
 * (graph_view.hpp:255-263).
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_graph b200_graph_t;

typedef struct b200_graph_view {
  /* POD mirror of edge_partition_device_view_t (edge_partition_device_view.cuh:355-483) plus
   * the host metadata of graph_view_t (graph_view.hpp:838+). Device pointers stay owned by
   * the graph. */
  int64_t num_vertices;
  int64_t num_edges;
  int32_t vertex_type;      /* B200_INT32 | B200_INT64 (ids as given by the caller)            */
  int32_t edge_type;        /* type of offsets: INT32 unless num_edges >= 2^31-1                */
  int32_t weight_type;      /* B200_FLOAT32 | B200_FLOAT64; meaningful when weights != NULL     */
  int32_t store_transposed; /* 1: rows are destinations (CSC, the pull layout)                  */
  int32_t renumbered;
  int32_t num_segments;     /* entries in segment_offsets minus 1 (4 for single GPU), 0 if none */
  int64_t segment_offsets[8]; /* [0, n_high, n_high+n_mid, n_nonzero, V] when renumbered        */
  const void* offsets;      /* edge_type  [num_vertices + 1]                                    */
  const void* indices;      /* int32      [num_edges]  (internal ids are always 32-bit)         */
  const void* weights;      /* weight_type[num_edges] or NULL                                   */
  const void* renumber_map; /* vertex_type[num_vertices]: internal id -> external id            */
  /* multi-GPU only (zero / NULL on a single GPU) */
  int32_t comm_rank, comm_size, row_comm_size, col_comm_size;
  int64_t local_vertex_first, local_vertex_last; /* this rank's vertex partition (internal ids) */
  /* ranges of this edge partition (graph_view.hpp:181-212); [0, V) on a single GPU.  `indices`
   * hold minor OFFSETS (global minor id - minor_first), the form the kernel gathers with. */
  int64_t major_first, major_last, minor_first, minor_last;
  /* DCSC form of the hypersparse segment of a multi-GPU edge partition
   * (compress_hypersparse_offsets, cpp/src/structure/detail/structure_utils.cuh:109-159;
   * edge_partition_device_view_t::dcs_nzd_vertices, edge_partition_device_view.cuh:159-351):
   * majors at or after major_hypersparse_first keep an offset only when they have edges.
   * dcs_offsets = offsets[0 .. hs] ++ offsets[r + 1] of every non-empty row r >= hs
   * (hs = major_hypersparse_first - major_first), dcs_nzd_vertices = those majors (global
   * internal ids).  NULL / 0 / -1 when the partition has no hypersparse segment (single GPU,
   * col_comm_size < 2).  `offsets` above stays the uncompressed array the kernels' layout is
   * built from. */
  const void* dcs_offsets;        /* edge_type [hs + dcs_nzd_count + 1]                          */
  const void* dcs_nzd_vertices;   /* int32     [dcs_nzd_count]                                   */
  int64_t dcs_nzd_count;
  int64_t major_hypersparse_first; /* -1 if none                                                  */
} b200_graph_view_t;

/* Single-GPU graph from a COO edge list (external ids).  vertices may be NULL (vertex set =
 * ids that appear in an edge, as cugraph_sg_graph_create does, c_api/graph_sg.cpp:61-190) or
 * list every vertex including isolated ones (as the C++ tests' Rmat_Usecase does,
 * cpp/tests/utilities/test_graphs.hpp:329-347).  renumber = 0 requires ids in [0, V). */
int b200_graph_create_sg(b200_handle_t* h, const void* src, const void* dst, const void* weights,
                         size_t num_edges, int vertex_type, int weight_type, const void* vertices,
                         size_t num_vertices, int store_transposed, int renumber,
                         b200_graph_t** out);
/* Multi-GPU graph: each rank passes the edges it currently holds (any distribution); they are
 * shuffled to their 2D owner (shuffle_wrappers.cu:41-174), vertices hash-assigned to ranks
 * (graph_utils.cuh:45-53), renumbered and compressed per local edge partition
 * (create_graph_from_edgelist_impl.cuh:207-555). */
int b200_graph_create_mg(b200_handle_t* h, const void* src, const void* dst, const void* weights,
                         size_t num_local_edges, int vertex_type, int weight_type,
                         const void* local_vertices, size_t num_local_vertices,
                         int store_transposed, b200_graph_t** out);
int b200_graph_view(const b200_graph_t* g, b200_graph_view_t* out);
/* view of local edge partition i (0 <= i < col_comm_size) of a multi-GPU graph */
int b200_graph_partition_view(const b200_graph_t* g, int i, b200_graph_view_t* out);
/* partition_t::vertex_partition_range_offsets (graph_view.hpp:121-170): out[comm_size + 1] */
int b200_graph_vertex_partition_offsets(const b200_graph_t* g, int64_t* out);
void b200_graph_free(b200_graph_t* g);

/* graph_view_t::compute_out_weight_sums (cpp/src/structure/graph_view_impl.cuh:238-265,
 * :696-721): per-source sum of edge weights (out-degree when unweighted), internal order.
 * out: value_type[local vertices]. Cached inside the graph after the first call. */
int b200_graph_out_weight_sums(b200_handle_t* h, b200_graph_t* g, int value_type, void* out);
/* in-degrees = per-row edge counts (graph_view_t::compute_in_degrees, graph_view_impl.cuh:203-219);
 * out: edge_type[local vertices] */
int b200_graph_in_degrees(b200_handle_t* h, const b200_graph_t* g, void* out);

/* -------------------------------------------------------------------------------------------
 * the hot path — per_v_transform_reduce_incoming_e
 * (cpp/src/prims/per_v_transform_reduce_incoming_outgoing_e.cuh:1082, :489-1038) specialised
 * to the edge ops of its three callers:  y[dst] = init + scale * sum_{src->dst} x[src] * w
 * x, y: value_type[num_vertices] in internal order (single GPU).
 * ------------------------------------------------------------------------------------------- */
int b200_spmv_pull(b200_handle_t* h, const b200_graph_t* g, int value_type, const void* x,
                   double scale, double init, void* y);

/* The same operation for callers that may not allocate (CUDA-graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Calls no cudaMalloc / cudaFree / allocating library routine and does not
 * synchronise: two kernels are enqueued on the handle's stream and the call returns.  Its scratch
 * (two partial sums per window of 4096 CSC positions) lives in a static __device__ array of
 * 4 MiB that every GPU loading the library holds until exit; graphs with more than 2^30 edges are
 * rejected with B200 error "invalid input" and a message naming the limit.  All calls share that
 * array: calls that overlap on one GPU (two streams, threads or handles) race - serialise them.
 * Single-GPU graphs; value_type B200_FLOAT32 / B200_FLOAT64; weights are read in their stored
 * type.  Same result as b200_spmv_pull up to floating-point association (sums are kept in fp64
 * and combined in a fixed order, so repeated calls are bit-identical). */
int b200_spmv_pull_noalloc(b200_handle_t* h, const b200_graph_t* g, int value_type, const void* x,
                           double scale, double init, void* y);

/* -------------------------------------------------------------------------------------------
 * drivers
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_centrality_stats {
  int64_t iterations;   /* power iterations executed                                           */
  int32_t converged;    /* 1 when the stop test passed                                         */
  double last_diff;     /* L1 difference of the last executed iteration                        */
  double elapsed_ms;    /* device time of the iteration loop (CUDA events on the handle stream) */
  int64_t gpu_launches; /* kernels launched by the loop                                        */
  double spmv_ms;       /* summed device time of the pull-kernel launches alone (CUDA events     */
                        /* around every launch of the first 256 iterations)                     */
  int64_t spmv_launches; /* launches covered by spmv_ms                                        */
} b200_centrality_stats_t;

typedef struct b200_pagerank_params {
  double alpha;
  double epsilon;
  uint64_t max_iterations;
  int32_t value_type;                  /* result_t: B200_FLOAT32 | B200_FLOAT64                 */
  int32_t has_initial_guess;           /* pageranks[] holds the guess on entry                  */
  const void* precomputed_out_weight_sums; /* value_type[local V], internal order, or NULL      */
  const void* personalization_vertices;    /* INTERNAL ids (int32) or NULL                      */
  const void* personalization_values;      /* value_type                                        */
  uint64_t personalization_size;
  int64_t fixed_iterations;            /* > 0: run exactly this many iterations, no stop test
                                          (parity at an identical iteration count; benchmarks)  */
} b200_pagerank_params_t;

/* cugraph::pagerank (cpp/include/cugraph/algorithms.hpp:1147-1159; driver
 * cpp/src/link_analysis/pagerank_impl.cuh:48-293).  pageranks: value_type[local V] (device),
 * internal order.  Returns CUGRAPH_UNKNOWN_ERROR with "PageRank failed to converge." exactly
 * where the reference throws (pagerank_impl.cuh:287-291). */
int b200_pagerank(b200_handle_t* h, b200_graph_t* g, const b200_pagerank_params_t* p,
                  void* pageranks, b200_centrality_stats_t* stats);

typedef struct b200_katz_params {
  double alpha;
  double beta;
  double epsilon;
  uint64_t max_iterations;
  int32_t value_type;
  int32_t has_initial_guess;
  int32_t normalize;
  const void* betas; /* value_type[local V] or NULL */
  int64_t fixed_iterations;
} b200_katz_params_t;

/* cugraph::katz_centrality (algorithms.hpp:1272-1283; katz_centrality_impl.cuh:42-173) */
int b200_katz_centrality(b200_handle_t* h, b200_graph_t* g, const b200_katz_params_t* p,
                         void* centralities, b200_centrality_stats_t* stats);

typedef struct b200_eigenvector_params {
  double epsilon;
  uint64_t max_iterations;
  int32_t value_type;
  int32_t has_initial_guess;
  int64_t fixed_iterations;
} b200_eigenvector_params_t;

/* cugraph::eigenvector_centrality (algorithms.hpp:1187; eigenvector_centrality_impl.cuh:45-138) */
int b200_eigenvector_centrality(b200_handle_t* h, b200_graph_t* g,
                                const b200_eigenvector_params_t* p, void* centralities,
                                b200_centrality_stats_t* stats);

typedef struct b200_hits_params {
  double epsilon;
  uint64_t max_iterations;
  int32_t value_type;
  int32_t has_initial_hubs_guess;  /* hubs[] holds the guess on entry (internal order)            */
  int32_t normalize;               /* 1: final scores are divided by their sums                    */
  int32_t pad;
  int64_t fixed_iterations;        /* > 0: run exactly this many iterations, no stop test          */
} b200_hits_params_t;

/* cugraph::hits (algorithms.hpp:1224-1234; hits_impl.cuh:54-185).  Single-GPU graphs stored
 * transposed.  stats->iterations = 0-based index of the iteration that met epsilon, or
 * max_iterations (the reference's final_iteration_count); stats->last_diff = L1 difference of
 * the hub scores of the last two iterations. */
int b200_hits(b200_handle_t* h, b200_graph_t* g, const b200_hits_params_t* p, void* hubs,
              void* authorities, b200_centrality_stats_t* stats);

/* -------------------------------------------------------------------------------------------
 * R-MAT generator — cugraph::generate_rmat_edgelist
 * (cpp/src/generators/generate_rmat_edgelist.cu:35-111).  Edges [first_edge, first_edge +
 * num_edges) of the stream identified by seed; ids are int32 (scale <= 31) or int64.
 * The random stream is this library's counter-based generator, not raft's (parity unpinned,
 * SURVEY.md 8c); the bit loop is the reference's.
 * ------------------------------------------------------------------------------------------- */
int b200_generate_rmat_edgelist(b200_handle_t* h, int scale, size_t num_edges, double a, double b,
                                double c, uint64_t seed, int clip_and_flip, uint64_t first_edge,
                                int vertex_type, void* src, void* dst);
/* uniform [0,1) edge weights from the same generator (test_graphs.hpp:279-284) */
int b200_uniform_fill(b200_handle_t* h, int value_type, void* out, size_t n, uint64_t seed);

/* -------------------------------------------------------------------------------------------
 * Edge-list ingest: the step before the path.  Host-side text -> COO in
 * host memory that is page-locked when a device is present (ready for one cudaMemcpyAsync into
 * b200_graph_create_sg).  No device work happens here.
 *
 * b200_read_matrix_market replaces read_edgelist_from_matrix_market_file / mm_properties /
 * mm_to_coo (cpp/tests/utilities/matrix_market_file_utilities.cu:53-222, :266-322): coordinate
 * files, pattern | real | integer values (complex and skew-symmetric are rejected as there,
 * :289), 1-based -> 0-based ids, and for symmetric files the mirrored copy of every off-diagonal
 * entry placed right after the entry (:199-215), so the edge order is the reference's.  Values
 * are read as double and narrowed to weight_type, pattern entries are 1.0 (:174-178).
 *
 * b200_read_csv_edgelist replaces the cudf.read_csv(path, delimiter=' ', dtype=['int32',
 * 'int32','float32'], header=None) call the reference's Python layer and datasets use
 * (python/cugraph/cugraph/testing/utils.py read_csv_file, experimental/datasets/dataset.py):
 * one edge per line, `src<delim>dst[<delim>weight]`, ids as written (no shift); the first
 * skip_rows lines are dropped; weight_type < 0 ignores a third column; a missing third column
 * with weight_type >= 0 is an error.
 *
 * The parse is split over `threads` host threads at line boundaries (0 = all hardware threads);
 * the result does not depend on the thread count.
 * ------------------------------------------------------------------------------------------- */
typedef struct b200_host_edgelist b200_host_edgelist_t;
typedef struct b200_host_edgelist_info {
  int64_t num_rows;      /* MTX size line (reference: number_of_vertices = rows); CSV: max id + 1 */
  int64_t num_cols;
  int64_t num_edges;     /* entries stored, mirrored copies included                            */
  int64_t file_entries;  /* entries in the file                                                 */
  int32_t is_symmetric;  /* MTX banner says symmetric (graph_properties_t::is_symmetric)         */
  int32_t is_pattern;    /* no values in the file                                               */
  int32_t vertex_type;   /* B200_INT32 | B200_INT64                                             */
  int32_t weight_type;   /* B200_FLOAT32 | B200_FLOAT64, or -1: no weights array                */
  int32_t pinned;        /* arrays are page-locked                                              */
  int32_t parser;        /* 0: parsed by the host reader, 1: by the device kernels              */
} b200_host_edgelist_info_t;
int b200_read_matrix_market(const char* path, int vertex_type, int weight_type, int threads,
                            b200_host_edgelist_t** out);
int b200_read_csv_edgelist(const char* path, char delimiter, int skip_rows, int vertex_type,
                           int weight_type, int threads, b200_host_edgelist_t** out);
int b200_host_edgelist_info(const b200_host_edgelist_t* e, b200_host_edgelist_info_t* out);
const void* b200_host_edgelist_src(const b200_host_edgelist_t* e);
const void* b200_host_edgelist_dst(const b200_host_edgelist_t* e);
const void* b200_host_edgelist_weights(const b200_host_edgelist_t* e); /* NULL if none */
void b200_host_edgelist_free(b200_host_edgelist_t* e);

/* The same two readers with the parse on the GPU: the text is copied to HBM once and parsed
 * there (one CTA per 65 KB tile staged through shared memory, one thread per 260-byte chunk;
 * count - scan - write), the COO arrays are left in device memory on h's device, ready for
 * b200_graph_create_sg.  Same files accepted, same arrays (order, ids, correctly rounded
 * values) and same errors as the host readers: whatever the kernels flag - a malformed line, a
 * line over 1 KB - is handed to the host reader.  info.pinned is 0. */
typedef struct b200_device_edgelist b200_device_edgelist_t;
int b200_read_matrix_market_device(b200_handle_t* h, const char* path, int vertex_type,
                                   int weight_type, b200_device_edgelist_t** out);
int b200_read_csv_edgelist_device(b200_handle_t* h, const char* path, char delimiter,
                                  int skip_rows, int vertex_type, int weight_type,
                                  b200_device_edgelist_t** out);
int b200_device_edgelist_info(const b200_device_edgelist_t* e, b200_host_edgelist_info_t* out);
const void* b200_device_edgelist_src(const b200_device_edgelist_t* e);
const void* b200_device_edgelist_dst(const b200_device_edgelist_t* e);
const void* b200_device_edgelist_weights(const b200_device_edgelist_t* e); /* NULL if none */
void b200_device_edgelist_free(b200_device_edgelist_t* e);

/* library identification: "sm_100a" build tag + version */
const char* b200_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_CUGRAPH_H */
