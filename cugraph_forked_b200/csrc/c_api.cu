// cugraph_c-compatible C API (include/cugraph_c/*.h) on top of the raw ABI: what
// cpp/src/c_api/{resource_handle,error,array,graph_sg,graph_mg,pagerank,katz,
// eigenvector_centrality,centrality_result}.cpp are for the reference.  Opaque handles are
// reinterpret_casts of the internal structs below, as in the reference (c_api/graph.hpp:31-44).
#include <cugraph_c/algorithms.h>

#include <memory>

#include "centrality.cuh"
#include "mg_comm.cuh"
#include "prims.cuh"

namespace b200 {
namespace capi {

struct error_t {
  std::string message;
};

struct resource_handle_t {
  b200_handle* h{nullptr};
  bool owned{false};
};

struct device_array_view_t {
  void* data{nullptr};
  size_t size{0};
  data_type_id_t type{INT32};
};

struct device_array_t {
  dbuf<char> buf;
  size_t size{0};
  data_type_id_t type{INT32};
  device_array_view_t self_view;  // cugraph_type_erased_device_array_view() returns a fresh copy
};

struct host_array_view_t {
  void* data{nullptr};
  size_t size{0};
  data_type_id_t type{INT32};
};

struct host_array_t {
  std::unique_ptr<char[]> buf;
  size_t size{0};
  data_type_id_t type{INT32};
};

struct graph_t {
  b200_graph* g{nullptr};
  data_type_id_t vertex_type{INT32};
  data_type_id_t weight_type{FLOAT32};
  bool multi_gpu{false};
};

struct centrality_result_t {
  device_array_t* vertex_ids{nullptr};
  device_array_t* values{nullptr};
};

// c_api/hits.cpp:34-40
struct hits_result_t {
  device_array_t* vertex_ids{nullptr};
  device_array_t* hubs{nullptr};
  device_array_t* authorities{nullptr};
  double hub_score_differences{0.0};
  size_t number_of_iterations{0};
};

inline size_t type_size(data_type_id_t t) { return (t == INT32 || t == FLOAT32) ? 4 : 8; }

inline cugraph_error_code_t fail(cugraph_error_t** error, int code, const char* msg)
{
  if (error) *error = reinterpret_cast<cugraph_error_t*>(new error_t{msg});
  return static_cast<cugraph_error_code_t>(code);
}

// CAPI_EXPECTS of the reference (c_api/error.hpp)
#define CAPI_EXPECTS(cond, code, msg, error) \
  do {                                       \
    if (!(cond)) return ::b200::capi::fail((error), (code), (msg)); \
  } while (0)

template <typename F>
cugraph_error_code_t run(cugraph_error_t** error, F&& f)
{
  try {
    f();
    return CUGRAPH_SUCCESS;
  } catch (b200::error const& e) {
    return fail(error, e.code, e.what());
  } catch (std::exception const& e) {
    return fail(error, kUnknownError, e.what());
  }
}

inline device_array_t* make_array(b200_handle* h, size_t n, data_type_id_t t)
{
  auto* a = new device_array_t();
  a->buf.alloc(n * type_size(t), h->stream);
  a->size = n;
  a->type = t;
  return a;
}

// scatter (vertex, value) pairs given with EXTERNAL ids into a dense local array
// (collect_local_vertex_values_from_ext_vertex_value_pairs, c_api/pagerank.cpp:156-210)
template <typename T>
__global__ void scatter_pairs_kernel(const int32_t* internal, const T* values, size_t n, T* dense,
                                     int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    int32_t v = internal[i];
    if (v < 0) { *bad = 1; continue; }
    dense[v] = values[i];
  }
}

// Multi-GPU: the (vertex, value) pairs a rank passes may belong to any rank; the reference
// shuffles them to the owner (shuffle_ext_vertex_value_pairs_to_local_gpu_by_vertex_partitioning,
// c_api/pagerank.cpp:125-155).  Here all pairs are all-gathered in rank order and every rank keeps
// the ones it owns, in that order (deterministic).  out_internal = local offsets, or global
// internal ids (local offset + local_first) when `global_ids`; an id no rank owns is an error.
template <typename T>
__global__ void keep_owned_kernel(const int32_t* internal, const uint32_t* pos, const T* values,
                                  size_t n, int32_t add, int32_t* out_internal, T* out_values)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    if (internal[i] < 0) continue;
    out_internal[pos[i]] = internal[i] + add;
    out_values[pos[i]]   = values[i];
  }
}
struct owned_flag { int32_t v; __host__ __device__ operator uint32_t() const { return v >= 0 ? 1u : 0u; } };

inline size_t owned_pairs(b200_handle* h, b200_graph* g, const device_array_view_t* vertices,
                          const device_array_view_t* values, data_type_id_t value_type,
                          bool global_ids, dbuf<int32_t>& out_internal, dbuf<char>& out_values)
{
  cudaStream_t s = h->stream;
  size_t n       = vertices->size;
  B200_EXPECTS(values->size == n, kInvalidInput, "vertex and value arrays differ in size");
  B200_EXPECTS(values->type == value_type, kInvalidInput,
               "value type must match the graph's weight type");
  const size_t vsz = type_size(vertices->type), esz = type_size(value_type);
  const void* vptr = vertices->data;
  const void* eptr = values->data;
  dbuf<char> all_v, all_e;
  size_t total = n;
  const bool mg = g->multi_gpu && h->size > 1;
  if (mg) {
    std::vector<int64_t> counts64 = allgather_host(h, h->world, h->size, {static_cast<int64_t>(n)});
    std::vector<size_t> counts(counts64.begin(), counts64.end());
    total = 0;
    for (size_t c : counts) total += c;
    all_v.alloc(std::max<size_t>(total, 1) * vsz, s);
    all_e.alloc(std::max<size_t>(total, 1) * esz, s);
    allgatherv(s, h->world, h->size, h->rank, vptr, all_v.data(), counts, vsz);
    allgatherv(s, h->world, h->size, h->rank, eptr, all_e.data(), counts, esz);
    vptr = all_v.data();
    eptr = all_e.data();
  }
  dbuf<int32_t> internal(std::max<size_t>(total, 1), s);
  dbuf<uint32_t> pos(total + 1, s);
  uint32_t kept = 0;
  if (total) {
    graph_ext_to_internal(h, g, vptr, total, internal.data());
    exclusive_scan<owned_flag, uint32_t>(reinterpret_cast<const owned_flag*>(internal.data()), pos.data(),
                                         total, s, true);
    CUDA_TRY(cudaMemcpyAsync(&kept, pos.data() + total, 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  int64_t kept_all = kept;
  if (mg) {
    kept_all = 0;
    for (int64_t k : allgather_host(h, h->world, h->size, {static_cast<int64_t>(kept)})) kept_all += k;
  }
  B200_EXPECTS(kept_all == static_cast<int64_t>(total), kInvalidInput,
               "Invalid input argument: vertex ids are not in the graph.");
  out_internal.alloc(std::max<size_t>(kept, 1), s);
  out_values.alloc(std::max<size_t>(kept, 1) * esz, s);
  if (total) {
    const int32_t add = global_ids ? static_cast<int32_t>(g->local_first) : 0;
    unsigned grid = static_cast<unsigned>(std::min<size_t>((total + 255) / 256, 4096));
    if (value_type == FLOAT32)
      keep_owned_kernel<float><<<grid, 256, 0, s>>>(internal.data(), pos.data(),
                                                   static_cast<const float*>(eptr), total, add,
                                                   out_internal.data(),
                                                   reinterpret_cast<float*>(out_values.data()));
    else
      keep_owned_kernel<double><<<grid, 256, 0, s>>>(internal.data(), pos.data(),
                                                    static_cast<const double*>(eptr), total, add,
                                                    out_internal.data(),
                                                    reinterpret_cast<double*>(out_values.data()));
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));   // the gathered inputs are freed on return
  }
  return kept;
}

inline void collect_dense(b200_handle* h, b200_graph* g, const device_array_view_t* vertices,
                          const device_array_view_t* values, data_type_id_t value_type,
                          dbuf<char>& dense)
{
  cudaStream_t s = h->stream;
  size_t nloc    = static_cast<size_t>(g->local_last - g->local_first);
  dbuf<int32_t> internal;
  dbuf<char> vals;
  const size_t n = owned_pairs(h, g, vertices, values, value_type, false, internal, vals);
  dense.alloc(nloc * type_size(value_type), s);
  CUDA_TRY(cudaMemsetAsync(dense.data(), 0, std::max<size_t>(nloc, 1) * type_size(value_type), s));
  dbuf<int> bad(1, s);
  CUDA_TRY(cudaMemsetAsync(bad.data(), 0, 4, s));
  if (n) {
    unsigned grid = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, 4096));
    if (value_type == FLOAT32)
      scatter_pairs_kernel<float><<<grid, 256, 0, s>>>(internal.data(),
                                                      reinterpret_cast<const float*>(vals.data()), n,
                                                      reinterpret_cast<float*>(dense.data()),
                                                      bad.data());
    else
      scatter_pairs_kernel<double><<<grid, 256, 0, s>>>(internal.data(),
                                                       reinterpret_cast<const double*>(vals.data()), n,
                                                       reinterpret_cast<double*>(dense.data()),
                                                       bad.data());
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
}

// result = { copy of the renumber map (external ids), scores in internal order }
inline centrality_result_t* make_result(b200_handle* h, graph_t* G, device_array_t* values)
{
  b200_graph* g  = G->g;
  size_t nloc    = static_cast<size_t>(g->local_last - g->local_first);
  auto* ids      = make_array(h, nloc, G->vertex_type);
  if (nloc)
    CUDA_TRY(cudaMemcpyAsync(ids->buf.data(), g->renumber_map.data(), nloc * type_size(G->vertex_type),
                             cudaMemcpyDeviceToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return new centrality_result_t{ids, values};
}

}  // namespace capi
}  // namespace b200

using namespace b200;
using namespace b200::capi;

extern "C" {

// ---- error -----------------------------------------------------------------------------
const char* cugraph_error_message(const cugraph_error_t* error)
{
  return error ? reinterpret_cast<const b200::capi::error_t*>(error)->message.c_str() : nullptr;
}
void cugraph_error_free(cugraph_error_t* error) { delete reinterpret_cast<b200::capi::error_t*>(error); }

// ---- resource handle ---------------------------------------------------------------------
cugraph_resource_handle_t* cugraph_create_resource_handle(void* raft_handle)
{
  auto* r = new resource_handle_t();
  if (raft_handle) {
    r->h = static_cast<b200_handle*>(raft_handle);
  } else {
    if (b200_handle_create(nullptr, &r->h) != 0) {
      delete r;
      return nullptr;
    }
    r->owned = true;
  }
  return reinterpret_cast<cugraph_resource_handle_t*>(r);
}
int cugraph_resource_handle_get_rank(const cugraph_resource_handle_t* handle)
{
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  return (r && r->h) ? r->h->rank : 0;
}
void cugraph_free_resource_handle(cugraph_resource_handle_t* handle)
{
  auto* r = reinterpret_cast<resource_handle_t*>(handle);
  if (!r) return;
  if (r->owned) b200_handle_free(r->h);
  delete r;
}

// ---- device arrays -------------------------------------------------------------------------
cugraph_error_code_t cugraph_type_erased_device_array_create(
  const cugraph_resource_handle_t* handle, size_t n_elems, data_type_id_t dtype,
  cugraph_type_erased_device_array_t** array, cugraph_error_t** error)
{
  *array = nullptr;
  *error = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(dtype >= INT32 && dtype < NTYPES, CUGRAPH_UNSUPPORTED_TYPE_COMBINATION,
               "invalid dtype", error);
  return run(error, [&] {
    *array = reinterpret_cast<cugraph_type_erased_device_array_t*>(make_array(r->h, n_elems, dtype));
  });
}

cugraph_error_code_t cugraph_type_erased_device_array_create_from_view(
  const cugraph_resource_handle_t* handle, const cugraph_type_erased_device_array_view_t* view,
  cugraph_type_erased_device_array_t** array, cugraph_error_t** error)
{
  *array = nullptr;
  *error = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* v = reinterpret_cast<const device_array_view_t*>(view);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(v != nullptr, CUGRAPH_INVALID_INPUT, "view is NULL", error);
  return run(error, [&] {
    auto* a = make_array(r->h, v->size, v->type);
    if (v->size)
      CUDA_TRY(cudaMemcpyAsync(a->buf.data(), v->data, v->size * type_size(v->type),
                               cudaMemcpyDeviceToDevice, r->h->stream));
    CUDA_TRY(cudaStreamSynchronize(r->h->stream));
    *array = reinterpret_cast<cugraph_type_erased_device_array_t*>(a);
  });
}

void cugraph_type_erased_device_array_free(cugraph_type_erased_device_array_t* p)
{
  delete reinterpret_cast<device_array_t*>(p);
}

cugraph_type_erased_device_array_view_t* cugraph_type_erased_device_array_view(
  cugraph_type_erased_device_array_t* array)
{
  auto* a = reinterpret_cast<device_array_t*>(array);
  return reinterpret_cast<cugraph_type_erased_device_array_view_t*>(
    new device_array_view_t{a->buf.data(), a->size, a->type});
}

cugraph_error_code_t cugraph_type_erased_device_array_view_as_type(
  cugraph_type_erased_device_array_t* array, data_type_id_t dtype,
  cugraph_type_erased_device_array_view_t** result_view, cugraph_error_t** error)
{
  *result_view = nullptr;
  *error       = nullptr;
  auto* a      = reinterpret_cast<device_array_t*>(array);
  CAPI_EXPECTS(a != nullptr, CUGRAPH_INVALID_INPUT, "array is NULL", error);
  // array.cpp: only same-width reinterpretation is allowed
  CAPI_EXPECTS(type_size(dtype) == type_size(a->type), CUGRAPH_INVALID_INPUT,
               "Could not treat type-erased array as the requested type: element sizes differ.",
               error);
  *result_view = reinterpret_cast<cugraph_type_erased_device_array_view_t*>(
    new device_array_view_t{a->buf.data(), a->size, dtype});
  return CUGRAPH_SUCCESS;
}

cugraph_type_erased_device_array_view_t* cugraph_type_erased_device_array_view_create(
  void* pointer, size_t n_elems, data_type_id_t dtype)
{
  return reinterpret_cast<cugraph_type_erased_device_array_view_t*>(
    new device_array_view_t{pointer, n_elems, dtype});
}
void cugraph_type_erased_device_array_view_free(cugraph_type_erased_device_array_view_t* p)
{
  delete reinterpret_cast<device_array_view_t*>(p);
}
size_t cugraph_type_erased_device_array_view_size(const cugraph_type_erased_device_array_view_t* p)
{
  return reinterpret_cast<const device_array_view_t*>(p)->size;
}
data_type_id_t cugraph_type_erased_device_array_view_type(
  const cugraph_type_erased_device_array_view_t* p)
{
  return reinterpret_cast<const device_array_view_t*>(p)->type;
}
const void* cugraph_type_erased_device_array_view_pointer(
  const cugraph_type_erased_device_array_view_t* p)
{
  return reinterpret_cast<const device_array_view_t*>(p)->data;
}

// ---- host arrays -------------------------------------------------------------------------
cugraph_error_code_t cugraph_type_erased_host_array_create(const cugraph_resource_handle_t* handle,
                                                           size_t n_elems, data_type_id_t dtype,
                                                           cugraph_type_erased_host_array_t** array,
                                                           cugraph_error_t** error)
{
  *array = nullptr;
  *error = nullptr;
  CAPI_EXPECTS(handle != nullptr, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  auto* a = new host_array_t();
  a->buf.reset(new char[std::max<size_t>(n_elems * type_size(dtype), 1)]);
  a->size = n_elems;
  a->type = dtype;
  *array  = reinterpret_cast<cugraph_type_erased_host_array_t*>(a);
  return CUGRAPH_SUCCESS;
}
void cugraph_type_erased_host_array_free(cugraph_type_erased_host_array_t* p)
{
  delete reinterpret_cast<host_array_t*>(p);
}
cugraph_type_erased_host_array_view_t* cugraph_type_erased_host_array_view(
  cugraph_type_erased_host_array_t* array)
{
  auto* a = reinterpret_cast<host_array_t*>(array);
  return reinterpret_cast<cugraph_type_erased_host_array_view_t*>(
    new host_array_view_t{a->buf.get(), a->size, a->type});
}
cugraph_type_erased_host_array_view_t* cugraph_type_erased_host_array_view_create(
  void* pointer, size_t n_elems, data_type_id_t dtype)
{
  return reinterpret_cast<cugraph_type_erased_host_array_view_t*>(
    new host_array_view_t{pointer, n_elems, dtype});
}
void cugraph_type_erased_host_array_view_free(cugraph_type_erased_host_array_view_t* p)
{
  delete reinterpret_cast<host_array_view_t*>(p);
}
size_t cugraph_type_erased_host_array_size(const cugraph_type_erased_host_array_view_t* p)
{
  return reinterpret_cast<const host_array_view_t*>(p)->size;
}
data_type_id_t cugraph_type_erased_host_array_type(const cugraph_type_erased_host_array_view_t* p)
{
  return reinterpret_cast<const host_array_view_t*>(p)->type;
}
void* cugraph_type_erased_host_array_pointer(const cugraph_type_erased_host_array_view_t* p)
{
  return reinterpret_cast<const host_array_view_t*>(p)->data;
}
cugraph_error_code_t cugraph_type_erased_host_array_view_copy(
  const cugraph_resource_handle_t*, cugraph_type_erased_host_array_view_t* dst,
  const cugraph_type_erased_host_array_view_t* src, cugraph_error_t** error)
{
  *error  = nullptr;
  auto* d = reinterpret_cast<host_array_view_t*>(dst);
  auto* s = reinterpret_cast<const host_array_view_t*>(src);
  CAPI_EXPECTS(d->size * type_size(d->type) == s->size * type_size(s->type), CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: dst and src differ in size.", error);
  std::memcpy(d->data, s->data, s->size * type_size(s->type));
  return CUGRAPH_SUCCESS;
}

// ---- copies ------------------------------------------------------------------------------
cugraph_error_code_t cugraph_type_erased_device_array_view_copy_from_host(
  const cugraph_resource_handle_t* handle, cugraph_type_erased_device_array_view_t* dst,
  const byte_t* h_src, cugraph_error_t** error)
{
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* d = reinterpret_cast<device_array_view_t*>(dst);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  return run(error, [&] {
    CUDA_TRY(cudaMemcpyAsync(d->data, h_src, d->size * type_size(d->type), cudaMemcpyHostToDevice,
                             r->h->stream));
    CUDA_TRY(cudaStreamSynchronize(r->h->stream));
  });
}
cugraph_error_code_t cugraph_type_erased_device_array_view_copy_to_host(
  const cugraph_resource_handle_t* handle, byte_t* h_dst,
  const cugraph_type_erased_device_array_view_t* src, cugraph_error_t** error)
{
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* s = reinterpret_cast<const device_array_view_t*>(src);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  return run(error, [&] {
    CUDA_TRY(cudaMemcpyAsync(h_dst, s->data, s->size * type_size(s->type), cudaMemcpyDeviceToHost,
                             r->h->stream));
    CUDA_TRY(cudaStreamSynchronize(r->h->stream));
  });
}
cugraph_error_code_t cugraph_type_erased_device_array_view_copy(
  const cugraph_resource_handle_t* handle, cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* src, cugraph_error_t** error)
{
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* d = reinterpret_cast<device_array_view_t*>(dst);
  auto* s = reinterpret_cast<const device_array_view_t*>(src);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(d->size * type_size(d->type) == s->size * type_size(s->type), CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: dst and src differ in size.", error);
  return run(error, [&] {
    CUDA_TRY(cudaMemcpyAsync(d->data, s->data, s->size * type_size(s->type),
                             cudaMemcpyDeviceToDevice, r->h->stream));
    CUDA_TRY(cudaStreamSynchronize(r->h->stream));
  });
}

// ---- graphs --------------------------------------------------------------------------------
static cugraph_error_code_t create_graph_common(
  const cugraph_resource_handle_t* handle, const cugraph_type_erased_device_array_view_t* src,
  const cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* weights,
  const cugraph_type_erased_device_array_view_t* edge_ids,
  const cugraph_type_erased_device_array_view_t* edge_types, bool_t store_transposed,
  bool_t renumber, bool multi_gpu, cugraph_graph_t** graph, cugraph_error_t** error)
{
  *graph = nullptr;
  *error = nullptr;
  auto* r  = reinterpret_cast<const resource_handle_t*>(handle);
  auto* ps = reinterpret_cast<const device_array_view_t*>(src);
  auto* pd = reinterpret_cast<const device_array_view_t*>(dst);
  auto* pw = reinterpret_cast<const device_array_view_t*>(weights);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  // the same checks, in the same order, as c_api/graph_sg.cpp:238-310
  CAPI_EXPECTS(ps->size == pd->size, CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: src size != dst size.", error);
  CAPI_EXPECTS(ps->type == pd->type, CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: src type != dst type.", error);
  CAPI_EXPECTS((weights == nullptr) || (pw->size == ps->size), CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: src size != weights size.", error);
  CAPI_EXPECTS((edge_types == nullptr && edge_ids == nullptr) ||
                 (edge_types != nullptr && edge_ids != nullptr),
               CUGRAPH_INVALID_INPUT,
               "Invalid input arguments: either none or both of edge ids and edge types must be "
               "provided.",
               error);
  CAPI_EXPECTS(ps->type == INT32 || ps->type == INT64, CUGRAPH_UNSUPPORTED_TYPE_COMBINATION,
               "vertex type must be INT32 or INT64", error);
  CAPI_EXPECTS(weights == nullptr || pw->type == FLOAT32 || pw->type == FLOAT64,
               CUGRAPH_UNSUPPORTED_TYPE_COMBINATION, "weight type must be FLOAT32 or FLOAT64", error);
  auto* G        = new graph_t();
  G->vertex_type = ps->type;
  G->weight_type = weights ? pw->type : FLOAT32;
  G->multi_gpu   = multi_gpu;
  int rc;
  if (multi_gpu)
    rc = b200_graph_create_mg(r->h, ps->data, pd->data, weights ? pw->data : nullptr, ps->size,
                              ps->type, G->weight_type, nullptr, 0, store_transposed, &G->g);
  else
    rc = b200_graph_create_sg(r->h, ps->data, pd->data, weights ? pw->data : nullptr, ps->size,
                              ps->type, G->weight_type, nullptr, 0, store_transposed, renumber,
                              &G->g);
  if (rc != 0) {
    delete G;
    return fail(error, rc, b200_last_error());
  }
  *graph = reinterpret_cast<cugraph_graph_t*>(G);
  return CUGRAPH_SUCCESS;
}

cugraph_error_code_t cugraph_sg_graph_create(
  const cugraph_resource_handle_t* handle, const cugraph_graph_properties_t*,
  const cugraph_type_erased_device_array_view_t* src,
  const cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* weights,
  const cugraph_type_erased_device_array_view_t* edge_ids,
  const cugraph_type_erased_device_array_view_t* edge_types, bool_t store_transposed,
  bool_t renumber, bool_t, cugraph_graph_t** graph, cugraph_error_t** error)
{
  return create_graph_common(handle, src, dst, weights, edge_ids, edge_types, store_transposed,
                             renumber, false, graph, error);
}

cugraph_error_code_t cugraph_mg_graph_create(
  const cugraph_resource_handle_t* handle, const cugraph_graph_properties_t*,
  const cugraph_type_erased_device_array_view_t* src,
  const cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* weights,
  const cugraph_type_erased_device_array_view_t* edge_ids,
  const cugraph_type_erased_device_array_view_t* edge_types, bool_t store_transposed, size_t,
  bool_t, cugraph_graph_t** graph, cugraph_error_t** error)
{
  // renumber is forced on for multi-GPU graphs (c_api/graph_mg.cpp:332)
  return create_graph_common(handle, src, dst, weights, edge_ids, edge_types, store_transposed,
                             TRUE, true, graph, error);
}

static void free_graph(cugraph_graph_t* graph)
{
  auto* G = reinterpret_cast<graph_t*>(graph);
  if (!G) return;
  b200_graph_free(G->g);
  delete G;
}
void cugraph_sg_graph_free(cugraph_graph_t* graph) { free_graph(graph); }
void cugraph_mg_graph_free(cugraph_graph_t* graph) { free_graph(graph); }

// ---- centrality results ----------------------------------------------------------------------
cugraph_type_erased_device_array_view_t* cugraph_centrality_result_get_vertices(
  cugraph_centrality_result_t* result)
{
  auto* r = reinterpret_cast<centrality_result_t*>(result);
  return cugraph_type_erased_device_array_view(
    reinterpret_cast<cugraph_type_erased_device_array_t*>(r->vertex_ids));
}
cugraph_type_erased_device_array_view_t* cugraph_centrality_result_get_values(
  cugraph_centrality_result_t* result)
{
  auto* r = reinterpret_cast<centrality_result_t*>(result);
  return cugraph_type_erased_device_array_view(
    reinterpret_cast<cugraph_type_erased_device_array_t*>(r->values));
}
void cugraph_centrality_result_free(cugraph_centrality_result_t* result)
{
  auto* r = reinterpret_cast<centrality_result_t*>(result);
  if (!r) return;
  delete r->vertex_ids;
  delete r->values;
  delete r;
}

// ---- algorithms ------------------------------------------------------------------------------
static cugraph_error_code_t pagerank_common(
  const cugraph_resource_handle_t* handle, cugraph_graph_t* graph,
  const cugraph_type_erased_device_array_view_t* ow_vertices,
  const cugraph_type_erased_device_array_view_t* ow_sums,
  const cugraph_type_erased_device_array_view_t* guess_vertices,
  const cugraph_type_erased_device_array_view_t* guess_values,
  const cugraph_type_erased_device_array_view_t* pers_vertices,
  const cugraph_type_erased_device_array_view_t* pers_values, double alpha, double epsilon,
  size_t max_iterations, cugraph_centrality_result_t** result, cugraph_error_t** error)
{
  *result = nullptr;
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* G = reinterpret_cast<graph_t*>(graph);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(G && G->g, CUGRAPH_INVALID_INPUT, "graph is NULL", error);
  auto view = [](const cugraph_type_erased_device_array_view_t* v) {
    return reinterpret_cast<const device_array_view_t*>(v);
  };
  // type checks of c_api/pagerank.cpp:256-303
  CAPI_EXPECTS(!ow_vertices || view(ow_vertices)->type == G->vertex_type, CUGRAPH_INVALID_INPUT,
               "vertex type of graph and precomputed_vertex_out_weight_vertices must match", error);
  CAPI_EXPECTS(!ow_sums || view(ow_sums)->type == G->weight_type, CUGRAPH_INVALID_INPUT,
               "vertex type of graph and precomputed_vertex_out_weight_sums must match", error);
  CAPI_EXPECTS(!guess_vertices || view(guess_vertices)->type == G->vertex_type,
               CUGRAPH_INVALID_INPUT,
               "vertex type of graph and initial_guess_vertices must match", error);
  CAPI_EXPECTS(!guess_values || view(guess_values)->type == G->weight_type, CUGRAPH_INVALID_INPUT,
               "vertex type of graph and initial_guess_values must match", error);
  CAPI_EXPECTS(!pers_vertices || view(pers_vertices)->type == G->vertex_type, CUGRAPH_INVALID_INPUT,
               "vertex type of graph and personalization_vector must match", error);
  CAPI_EXPECTS(!pers_values || view(pers_values)->type == G->weight_type, CUGRAPH_INVALID_INPUT,
               "vertex type of graph and personalization_vector must match", error);
  return run(error, [&] {
    b200_handle* h = r->h;
    b200_graph* g  = G->g;
    // PageRank expects store_transposed == true: transpose in place (c_api/pagerank.cpp:105-110)
    if (!g->store_transposed) graph_transpose_storage(h, g);
    size_t nloc = static_cast<size_t>(g->local_last - g->local_first);
    std::unique_ptr<device_array_t> values(make_array(h, nloc, G->weight_type));
    b200_pagerank_params_t p{};
    p.alpha          = alpha;
    p.epsilon        = epsilon;
    p.max_iterations = max_iterations;
    p.value_type     = G->weight_type;
    dbuf<char> ow_dense, guess_dense;
    dbuf<int32_t> pers_internal;
    if (ow_sums) {
      collect_dense(h, g, view(ow_vertices), view(ow_sums), G->weight_type, ow_dense);
      p.precomputed_out_weight_sums = ow_dense.data();
    }
    if (guess_values) {
      collect_dense(h, g, view(guess_vertices), view(guess_values), G->weight_type, guess_dense);
      CUDA_TRY(cudaMemcpyAsync(values->buf.data(), guess_dense.data(),
                               nloc * type_size(G->weight_type), cudaMemcpyDeviceToDevice, h->stream));
      p.has_initial_guess = 1;
    }
    dbuf<char> pers_owned;
    if (pers_vertices) {
      // the pairs this rank owns, as GLOBAL internal ids (the convention of b200_pagerank on a
      // multi-GPU graph; on a single GPU local_first is 0)
      const size_t np = owned_pairs(h, g, view(pers_vertices), view(pers_values), G->weight_type, true,
                                    pers_internal, pers_owned);
      p.personalization_vertices = pers_internal.data();
      p.personalization_values   = pers_owned.data();
      p.personalization_size     = np;
    }
    pagerank(h, g, &p, values->buf.data(), nullptr);
    *result = reinterpret_cast<cugraph_centrality_result_t*>(make_result(h, G, values.release()));
  });
}

cugraph_error_code_t cugraph_pagerank(
  const cugraph_resource_handle_t* handle, cugraph_graph_t* graph,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_vertices,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_sums,
  const cugraph_type_erased_device_array_view_t* initial_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_guess_values, double alpha, double epsilon,
  size_t max_iterations, bool_t, cugraph_centrality_result_t** result, cugraph_error_t** error)
{
  return pagerank_common(handle, graph, precomputed_vertex_out_weight_vertices,
                         precomputed_vertex_out_weight_sums, initial_guess_vertices,
                         initial_guess_values, nullptr, nullptr, alpha, epsilon, max_iterations,
                         result, error);
}

cugraph_error_code_t cugraph_personalized_pagerank(
  const cugraph_resource_handle_t* handle, cugraph_graph_t* graph,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_vertices,
  const cugraph_type_erased_device_array_view_t* precomputed_vertex_out_weight_sums,
  const cugraph_type_erased_device_array_view_t* initial_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_guess_values,
  const cugraph_type_erased_device_array_view_t* personalization_vertices,
  const cugraph_type_erased_device_array_view_t* personalization_values, double alpha,
  double epsilon, size_t max_iterations, bool_t, cugraph_centrality_result_t** result,
  cugraph_error_t** error)
{
  return pagerank_common(handle, graph, precomputed_vertex_out_weight_vertices,
                         precomputed_vertex_out_weight_sums, initial_guess_vertices,
                         initial_guess_values, personalization_vertices, personalization_values,
                         alpha, epsilon, max_iterations, result, error);
}

cugraph_error_code_t cugraph_katz_centrality(const cugraph_resource_handle_t* handle,
                                             cugraph_graph_t* graph,
                                             const cugraph_type_erased_device_array_view_t*,
                                             double alpha, double beta, double epsilon,
                                             size_t max_iterations, bool_t,
                                             cugraph_centrality_result_t** result,
                                             cugraph_error_t** error)
{
  *result = nullptr;
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* G = reinterpret_cast<graph_t*>(graph);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(G && G->g, CUGRAPH_INVALID_INPUT, "graph is NULL", error);
  return run(error, [&] {
    b200_handle* h = r->h;
    b200_graph* g  = G->g;
    if (!g->store_transposed) graph_transpose_storage(h, g);
    size_t nloc = static_cast<size_t>(g->local_last - g->local_first);
    std::unique_ptr<device_array_t> values(make_array(h, nloc, G->weight_type));
    // as the reference: betas are NOT forwarded (c_api/katz.cpp:158-159 passes nullptr),
    // has_initial_guess = false, normalize = true (katz.cpp:129-130)
    b200_katz_params_t p{};
    p.alpha          = alpha;
    p.beta           = beta;
    p.epsilon        = epsilon;
    p.max_iterations = max_iterations;
    p.value_type     = G->weight_type;
    p.normalize      = 1;
    katz_centrality(h, g, &p, values->buf.data(), nullptr);
    *result = reinterpret_cast<cugraph_centrality_result_t*>(make_result(h, G, values.release()));
  });
}

cugraph_error_code_t cugraph_eigenvector_centrality(const cugraph_resource_handle_t* handle,
                                                    cugraph_graph_t* graph, double epsilon,
                                                    size_t max_iterations, bool_t,
                                                    cugraph_centrality_result_t** result,
                                                    cugraph_error_t** error)
{
  *result = nullptr;
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* G = reinterpret_cast<graph_t*>(graph);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(G && G->g, CUGRAPH_INVALID_INPUT, "graph is NULL", error);
  return run(error, [&] {
    b200_handle* h = r->h;
    b200_graph* g  = G->g;
    if (!g->store_transposed) graph_transpose_storage(h, g);
    size_t nloc = static_cast<size_t>(g->local_last - g->local_first);
    std::unique_ptr<device_array_t> values(make_array(h, nloc, G->weight_type));
    b200_eigenvector_params_t p{};
    p.epsilon        = epsilon;
    p.max_iterations = max_iterations;
    p.value_type     = G->weight_type;
    eigenvector_centrality(h, g, &p, values->buf.data(), nullptr);
    *result = reinterpret_cast<cugraph_centrality_result_t*>(make_result(h, G, values.release()));
  });
}

// ---- HITS (c_api/hits.cpp) --------------------------------------------------------------------
cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_vertices(cugraph_hits_result_t* result)
{
  auto* r = reinterpret_cast<hits_result_t*>(result);
  return cugraph_type_erased_device_array_view(
    reinterpret_cast<cugraph_type_erased_device_array_t*>(r->vertex_ids));
}
cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_hubs(cugraph_hits_result_t* result)
{
  auto* r = reinterpret_cast<hits_result_t*>(result);
  return cugraph_type_erased_device_array_view(
    reinterpret_cast<cugraph_type_erased_device_array_t*>(r->hubs));
}
cugraph_type_erased_device_array_view_t* cugraph_hits_result_get_authorities(
  cugraph_hits_result_t* result)
{
  auto* r = reinterpret_cast<hits_result_t*>(result);
  return cugraph_type_erased_device_array_view(
    reinterpret_cast<cugraph_type_erased_device_array_t*>(r->authorities));
}
double cugraph_hits_result_get_hub_score_differences(cugraph_hits_result_t* result)
{
  return reinterpret_cast<hits_result_t*>(result)->hub_score_differences;
}
size_t cugraph_hits_result_get_number_of_iterations(cugraph_hits_result_t* result)
{
  return reinterpret_cast<hits_result_t*>(result)->number_of_iterations;
}
void cugraph_hits_result_free(cugraph_hits_result_t* result)
{
  auto* r = reinterpret_cast<hits_result_t*>(result);
  if (!r) return;
  delete r->vertex_ids;
  delete r->hubs;
  delete r->authorities;
  delete r;
}

cugraph_error_code_t cugraph_hits(
  const cugraph_resource_handle_t* handle, cugraph_graph_t* graph, double epsilon,
  size_t max_iterations, const cugraph_type_erased_device_array_view_t* initial_hubs_guess_vertices,
  const cugraph_type_erased_device_array_view_t* initial_hubs_guess_values, bool_t normalize, bool_t,
  cugraph_hits_result_t** result, cugraph_error_t** error)
{
  *result = nullptr;
  *error  = nullptr;
  auto* r = reinterpret_cast<const resource_handle_t*>(handle);
  auto* G = reinterpret_cast<graph_t*>(graph);
  CAPI_EXPECTS(r && r->h, CUGRAPH_INVALID_HANDLE, "invalid resource handle", error);
  CAPI_EXPECTS(G && G->g, CUGRAPH_INVALID_INPUT, "graph is NULL", error);
  auto view = [](const cugraph_type_erased_device_array_view_t* v) {
    return reinterpret_cast<const device_array_view_t*>(v);
  };
  // hits.cpp:228-239
  CAPI_EXPECTS((initial_hubs_guess_vertices == nullptr) == (initial_hubs_guess_values == nullptr),
               CUGRAPH_INVALID_INPUT,
               "initial_hubs_guess_vertices and initial_hubs_guess_values must be given together",
               error);
  CAPI_EXPECTS(!initial_hubs_guess_vertices ||
                 view(initial_hubs_guess_vertices)->type == G->vertex_type,
               CUGRAPH_INVALID_INPUT,
               "vertex type of graph and initial_hubs_guess_vertices must match", error);
  CAPI_EXPECTS(!initial_hubs_guess_values || view(initial_hubs_guess_values)->type == G->weight_type,
               CUGRAPH_INVALID_INPUT,
               "weight type of graph and initial_hubs_guess_values must match", error);
  return run(error, [&] {
    b200_handle* h = r->h;
    b200_graph* g  = G->g;
    // HITS expects store_transposed == true (hits.cpp:94-100)
    if (!g->store_transposed) graph_transpose_storage(h, g);
    size_t nloc = static_cast<size_t>(g->local_last - g->local_first);
    std::unique_ptr<device_array_t> hubs(make_array(h, nloc, G->weight_type));
    std::unique_ptr<device_array_t> auth(make_array(h, nloc, G->weight_type));
    b200_hits_params_t p{};
    p.epsilon        = epsilon;
    p.max_iterations = max_iterations;
    p.value_type     = G->weight_type;
    p.normalize      = normalize ? 1 : 0;
    dbuf<char> guess_dense;
    if (initial_hubs_guess_values) {
      collect_dense(h, g, view(initial_hubs_guess_vertices), view(initial_hubs_guess_values),
                    G->weight_type, guess_dense);
      CUDA_TRY(cudaMemcpyAsync(hubs->buf.data(), guess_dense.data(), nloc * type_size(G->weight_type),
                               cudaMemcpyDeviceToDevice, h->stream));
      p.has_initial_hubs_guess = 1;
    }
    b200_centrality_stats_t st{};
    hits(h, g, &p, hubs->buf.data(), auth->buf.data(), &st);
    std::unique_ptr<centrality_result_t> ids(make_result(h, G, nullptr));
    auto* out = new hits_result_t{ids->vertex_ids, hubs.release(), auth.release(), st.last_diff,
                                  static_cast<size_t>(st.iterations)};
    *result = reinterpret_cast<cugraph_hits_result_t*>(out);
  });
}

}  // extern "C"
