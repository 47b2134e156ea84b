// extern "C" surface of libcugraph_b200.so (include/b200_cugraph.h): argument checks, exception
// -> error-code translation (what cpp/src/c_api/utils.hpp:25-56 `run_algorithm` does for the
// reference) and the handle object.
#include <cstdlib>
#include <cstring>

#include "centrality.cuh"

namespace b200 {
static thread_local std::string g_last_error;
void set_last_error(const char* msg) { g_last_error = msg ? msg : ""; }

template <typename F>
static int guarded(F&& f)
{
  try {
    f();
    return kSuccess;
  } catch (error const& e) {
    set_last_error(e.what());
    return e.code;
  } catch (std::exception const& e) {
    set_last_error(e.what());
    return kUnknownError;
  } catch (...) {
    set_last_error("unknown exception");
    return kUnknownError;
  }
}

static void fill_device_info(b200_handle* h)
{
  CUDA_TRY(cudaGetDevice(&h->device));
  cudaDeviceProp prop{};
  CUDA_TRY(cudaGetDeviceProperties(&prop, h->device));
  h->sm_count          = prop.multiProcessorCount;
  h->l2_bytes          = static_cast<size_t>(prop.l2CacheSize);
  // random 4/8-byte gathers dominate: a 32-byte L2 fetch granularity halves the DRAM bytes of a
  // gather that misses (the default fetches 64); B200_L2_FETCH=64|128 restores / widens it
  {
    const char* e = std::getenv("B200_L2_FETCH");
    const size_t g = e ? static_cast<size_t>(std::atoll(e)) : 32;
    if (g == 32 || g == 64 || g == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g);
    cudaGetLastError();
  }
  B200_EXPECTS(prop.major >= 10, kUnknownError,
               "libcugraph_b200 is built for sm_100a only: no Blackwell GPU on this device.");
  // keep freed blocks in the default pool (dbuf allocates from it, common.cuh); the
  // B200_POOL_RELEASE_MB environment variable caps what is kept
  cudaMemPool_t pool{};
  CUDA_TRY(cudaDeviceGetDefaultMemPool(&pool, h->device));
  unsigned long long keep = ~0ull;
  if (const char* v = std::getenv("B200_POOL_RELEASE_MB")) keep = std::strtoull(v, nullptr, 10) << 20;
  CUDA_TRY(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
}
}  // namespace b200

namespace {
template <typename T>
__global__ void from_double_kernel(const double* in, size_t n, T* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<T>(in[i]);
}
}  // namespace

using namespace b200;

extern "C" {

const char* b200_last_error(void) { return g_last_error.c_str(); }

const char* b200_build_info(void) { return "libcugraph_b200 0.2 (sm_100a, CUDA " B200_STR(CUDART_VERSION) "; spmv sell-r2a)"; }

int b200_handle_create(void* stream, b200_handle_t** out)
{
  return guarded([&] {
    B200_EXPECTS(out != nullptr, kInvalidInput, "out == NULL");
    *out      = nullptr;
    auto* h   = new b200_handle();
    try {
      fill_device_info(h);
      if (stream) {
        h->stream = static_cast<cudaStream_t>(stream);
      } else {
        CUDA_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
      }
    } catch (...) {
      delete h;
      throw;
    }
    *out = h;
  });
}

int b200_nccl_unique_id(void* out_128_bytes)
{
  return guarded([&] {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId id;
    NCCL_TRY(ncclGetUniqueId(&id));
    std::memcpy(out_128_bytes, &id, sizeof id);
  });
}

int b200_handle_create_mg(void* stream, const void* nccl_unique_id, int rank, int world_size,
                          int row_comm_size, b200_handle_t** out)
{
  return guarded([&] {
    B200_EXPECTS(out != nullptr && nccl_unique_id != nullptr, kInvalidInput, "NULL argument");
    B200_EXPECTS(world_size >= 1 && rank >= 0 && rank < world_size, kInvalidInput, "bad rank");
    B200_EXPECTS(row_comm_size >= 1 && world_size % row_comm_size == 0, kInvalidInput,
                 "row_comm_size must divide the world size");
    *out    = nullptr;
    auto* h = new b200_handle();
    try {
      fill_device_info(h);
      if (stream) {
        h->stream = static_cast<cudaStream_t>(stream);
      } else {
        CUDA_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        h->own_stream = true;
      }
      ncclUniqueId id;
      std::memcpy(&id, nccl_unique_id, sizeof id);
      NCCL_TRY(ncclCommInitRank(&h->world, world_size, id, rank));
      h->rank     = rank;
      h->size     = world_size;
      h->row_size = row_comm_size;
      h->col_size = world_size / row_comm_size;
      // partition_manager.hpp:86-103: row_comm = ranks sharing rank / R, col_comm = ranks
      // sharing rank % R
      int row_color = rank / row_comm_size;
      int col_color = rank % row_comm_size;
      NCCL_TRY(ncclCommSplit(h->world, row_color, rank, &h->row, nullptr));
      NCCL_TRY(ncclCommSplit(h->world, col_color, rank, &h->col, nullptr));
      h->row_rank = rank % row_comm_size;
      h->col_rank = rank / row_comm_size;
      CUDA_TRY(cudaStreamCreateWithFlags(&h->comm_stream, cudaStreamNonBlocking));
    } catch (...) {
      delete h;
      throw;
    }
    *out = h;
  });
}

int b200_handle_rank(const b200_handle_t* h) { return h ? h->rank : 0; }
int b200_handle_world_size(const b200_handle_t* h) { return h ? h->size : 1; }

void b200_handle_free(b200_handle_t* h)
{
  if (!h) return;
  if (h->row) ncclCommDestroy(h->row);
  if (h->col) ncclCommDestroy(h->col);
  if (h->world) ncclCommDestroy(h->world);
  if (h->comm_stream) cudaStreamDestroy(h->comm_stream);
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

int b200_graph_create_sg(b200_handle_t* h, const void* src, const void* dst, const void* weights,
                         size_t num_edges, int vertex_type, int weight_type, const void* vertices,
                         size_t num_vertices, int store_transposed, int renumber,
                         b200_graph_t** out)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(out != nullptr, kInvalidInput, "out == NULL");
    *out = nullptr;
    B200_EXPECTS(vertex_type == B200_INT32 || vertex_type == B200_INT64,
                 kUnsupportedTypeCombination, "vertex type must be INT32 or INT64");
    B200_EXPECTS(!weights || weight_type == B200_FLOAT32 || weight_type == B200_FLOAT64,
                 kUnsupportedTypeCombination, "weight type must be FLOAT32 or FLOAT64");
    B200_EXPECTS(num_edges == 0 || (src && dst), kInvalidInput, "src / dst == NULL");
    auto* g = new b200_graph();
    try {
      graph_create_sg(h, src, dst, weights, num_edges, vertex_type, weight_type, vertices,
                      num_vertices, store_transposed != 0, renumber != 0, g);
    } catch (...) {
      delete g;
      throw;
    }
    *out = g;
  });
}

int b200_graph_create_mg(b200_handle_t* h, const void* src, const void* dst, const void* weights,
                         size_t num_local_edges, int vertex_type, int weight_type,
                         const void* local_vertices, size_t num_local_vertices,
                         int store_transposed, b200_graph_t** out)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr && h->world != nullptr, kInvalidHandle,
                 "multi-GPU graph needs a handle from b200_handle_create_mg");
    B200_EXPECTS(out != nullptr, kInvalidInput, "out == NULL");
    *out    = nullptr;
    auto* g = new b200_graph();
    try {
      graph_create_mg(h, src, dst, weights, num_local_edges, vertex_type, weight_type,
                      local_vertices, num_local_vertices, store_transposed != 0, g);
    } catch (...) {
      delete g;
      throw;
    }
    *out = g;
  });
}

static void fill_view(const b200_graph* g, int part, b200_graph_view_t* v)
{
  std::memset(v, 0, sizeof *v);
  const edge_partition& ep = g->parts[part];
  v->num_vertices          = g->num_vertices;
  v->num_edges             = g->multi_gpu ? ep.num_edges : g->num_edges;
  v->vertex_type           = g->vertex_type;
  v->edge_type             = ep.offsets64 ? B200_INT64 : B200_INT32;
  v->weight_type           = g->weight_type;
  v->store_transposed      = g->store_transposed;
  v->renumbered            = g->renumbered;
  v->num_segments          = g->num_segments;
  for (int i = 0; i < 8; ++i) v->segment_offsets[i] = g->segment_offsets[i];
  v->offsets      = ep.offsets.data();
  v->indices      = ep.indices.data();
  v->weights      = g->weighted ? ep.weights.data() : nullptr;
  v->renumber_map = g->renumber_map.data();
  v->local_vertex_first = g->local_first;
  v->local_vertex_last  = g->local_last;
  v->major_first        = ep.major_first;
  v->major_last         = ep.major_first + ep.num_rows;
  v->minor_first        = ep.minor_first;
  v->minor_last         = ep.minor_first + ep.num_minors;
  v->major_hypersparse_first = -1;
  if (ep.hypersparse_first >= 0) {
    v->dcs_offsets             = ep.dcs_offsets.data();
    v->dcs_nzd_vertices        = ep.dcs_nzd_vertices.data();
    v->dcs_nzd_count           = ep.dcs_nzd_count;
    v->major_hypersparse_first = ep.major_first + ep.hypersparse_first;
  }
  if (g->multi_gpu && part < static_cast<int>(g->partition_segment_offsets.size())) {
    auto const& seg = g->partition_segment_offsets[part];
    v->num_segments = static_cast<int32_t>(seg.size()) - 1;
    for (size_t i = 0; i < 8; ++i) v->segment_offsets[i] = i < seg.size() ? seg[i] : 0;
  }
}

int b200_graph_view(const b200_graph_t* g, b200_graph_view_t* out)
{
  return guarded([&] {
    B200_EXPECTS(g && out, kInvalidInput, "NULL argument");
    fill_view(g, 0, out);
    if (g->multi_gpu) {
      // the whole-graph view of an MG graph: this rank's vertex partition and its segments
      out->num_edges    = g->num_edges;
      out->num_segments = g->num_segments;
      for (int i = 0; i < 8; ++i) out->segment_offsets[i] = g->segment_offsets[i];
    }
  });
}

int b200_graph_vertex_partition_offsets(const b200_graph_t* g, int64_t* out)
{
  return guarded([&] {
    B200_EXPECTS(g && out, kInvalidInput, "NULL argument");
    if (g->multi_gpu) {
      for (size_t i = 0; i < g->vertex_partition_offsets.size(); ++i)
        out[i] = g->vertex_partition_offsets[i];
    } else {
      out[0] = 0;
      out[1] = g->num_vertices;
    }
  });
}

int b200_graph_partition_view(const b200_graph_t* g, int i, b200_graph_view_t* out)
{
  return guarded([&] {
    B200_EXPECTS(g && out, kInvalidInput, "NULL argument");
    B200_EXPECTS(i >= 0 && i < static_cast<int>(g->parts.size()), kInvalidInput,
                 "partition index out of range");
    fill_view(g, i, out);
    out->num_vertices = g->parts[i].num_rows;
  });
}

void b200_graph_free(b200_graph_t* g) { delete g; }


int b200_graph_out_weight_sums(b200_handle_t* h, b200_graph_t* g, int value_type, void* out)
{
  return guarded([&] {
    B200_EXPECTS(h && g && out, kInvalidInput, "NULL argument");
    graph_out_weight_sums(h, g);
    size_t n = static_cast<size_t>(g->local_last - g->local_first);
    if (n == 0) return;
    unsigned grid = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, 4096));
    if (value_type == B200_FLOAT32)
      from_double_kernel<float><<<grid, 256, 0, h->stream>>>(g->out_weight_sums.data(), n,
                                                            static_cast<float*>(out));
    else
      from_double_kernel<double><<<grid, 256, 0, h->stream>>>(g->out_weight_sums.data(), n,
                                                             static_cast<double*>(out));
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(h->stream));
  });
}

int b200_graph_in_degrees(b200_handle_t* h, const b200_graph_t* g, void* out)
{
  return guarded([&] {
    B200_EXPECTS(h && g && out, kInvalidInput, "NULL argument");
    B200_EXPECTS(!g->multi_gpu, kNotImplemented, "in-degrees of a multi-GPU graph");
    graph_in_degrees(h, g, out);
  });
}

int b200_spmv_pull(b200_handle_t* h, const b200_graph_t* g, int value_type, const void* x,
                   double scale, double init, void* y)
{
  return guarded([&] {
    B200_EXPECTS(h && g && x && y, kInvalidInput, "NULL argument");
    spmv_pull(h, const_cast<b200_graph_t*>(g), value_type, x, scale, init, y);
  });
}

int b200_pagerank(b200_handle_t* h, b200_graph_t* g, const b200_pagerank_params_t* p,
                  void* pageranks, b200_centrality_stats_t* stats)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(g && p && (pageranks || g->local_last == g->local_first), kInvalidInput,
                 "NULL argument");
    pagerank(h, g, p, pageranks, stats);
  });
}

int b200_katz_centrality(b200_handle_t* h, b200_graph_t* g, const b200_katz_params_t* p,
                         void* centralities, b200_centrality_stats_t* stats)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(g && p, kInvalidInput, "NULL argument");
    katz_centrality(h, g, p, centralities, stats);
  });
}

int b200_eigenvector_centrality(b200_handle_t* h, b200_graph_t* g,
                                const b200_eigenvector_params_t* p, void* centralities,
                                b200_centrality_stats_t* stats)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(g && p, kInvalidInput, "NULL argument");
    eigenvector_centrality(h, g, p, centralities, stats);
  });
}

int b200_hits(b200_handle_t* h, b200_graph_t* g, const b200_hits_params_t* p, void* hubs,
              void* authorities, b200_centrality_stats_t* stats)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(g && p, kInvalidInput, "NULL argument");
    hits(h, g, p, hubs, authorities, stats);
  });
}

int b200_generate_rmat_edgelist(b200_handle_t* h, int scale, size_t num_edges, double a, double b,
                                double c, uint64_t seed, int clip_and_flip, uint64_t first_edge,
                                int vertex_type, void* src, void* dst)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    generate_rmat(h, scale, num_edges, a, b, c, seed, clip_and_flip != 0, first_edge, vertex_type,
                  src, dst);
  });
}

int b200_uniform_fill(b200_handle_t* h, int value_type, void* out, size_t n, uint64_t seed)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    uniform_fill(h, value_type, out, n, seed);
  });
}

}  // extern "C"
