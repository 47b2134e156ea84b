// HITS on top of the pull SpMV (the first row after the hot path).
//   cpp/src/link_analysis/hits_impl.cuh:54-185   (driver: authorities = pull over the CSC,
//   hubs = per_v_transform_reduce_outgoing_e, max-normalisation of both, L1 stop test on hubs)
//   cpp/include/cugraph/algorithms.hpp:1224-1234 (cugraph::hits signature)
// The reference runs the second half as a scatter with atomics over the same CSC
// (per_v_transform_reduce_incoming_outgoing_e.cuh:154-178, :972-1037).  Here both halves are
// pulls: hubs[s] = sum over the out-edges of s of authorities[dst] is a pull over the CSR of
// the same graph, which is built once per graph (same vertex numbering, no weights) and cached;
// no atomics, so the result is deterministic.  Weights play no role in HITS; a weighted graph
// gets a weight-less CSC companion as well.  Single GPU only in this round.
#include <cmath>

#include "centrality.cuh"

namespace b200 {

namespace {

constexpr int kHT = 256;
inline unsigned hgrid(size_t n, int sm_count)
{
  size_t b   = (n + kHT * 4 - 1) / (kHT * 4);
  size_t cap = static_cast<size_t>(sm_count) * 4;
  if (b > cap) b = cap;
  if (b == 0) b = 1;
  return static_cast<unsigned>(b);
}

// majors[i] = row of entry i, minors[i] = indices[i]   (one warp per row)
template <typename ET>
__global__ void expand_rows_kernel(const ET* offsets, const int32_t* indices, int64_t num_rows,
                                   int32_t* majors, int32_t* minors)
{
  const int lane     = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nw   = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < num_rows; r += nw) {
    const int64_t b = offsets[r], e = offsets[r + 1];
    for (int64_t i = b + lane; i < e; i += 32) {
      majors[i] = static_cast<int32_t>(r);
      minors[i] = indices[i];
    }
  }
}

__global__ void iota32_kernel(int32_t* out, size_t n)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<int32_t>(i);
}

template <typename RT>
__global__ void fill_kernel(RT* out, size_t n, RT v)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = v;
}

// MODE 0: max(a)   MODE 1: sum(a)   MODE 2: sum |a - b|     -> partials[blockIdx.x] (double)
template <typename RT, int MODE>
__global__ void __launch_bounds__(kHT) reduce_partials_kernel(const RT* a, const RT* b, size_t n,
                                                              double* partials)
{
  __shared__ double scratch[32];
  double acc = 0.0;   // all three reductions run over non-negative values
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const double v = static_cast<double>(a[i]);
    if (MODE == 0) acc = v > acc ? v : acc;
    else if (MODE == 1) acc += v;
    else acc += fabs(v - static_cast<double>(b[i]));
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (MODE == 0) acc = warp_max(acc); else acc = warp_sum(acc);
  if (lane == 0) scratch[w] = acc;
  __syncthreads();
  if (w == 0) {
    double r = lane < kHT / 32 ? scratch[lane] : 0.0;
    if (MODE == 0) r = warp_max(r); else r = warp_sum(r);
    if (lane == 0) partials[blockIdx.x] = r;
  }
}

// one block folds the partials in a fixed order
template <int MODE>
__global__ void __launch_bounds__(kHT) fold_partials_kernel(const double* partials, int n,
                                                            double* out)
{
  __shared__ double scratch[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += kHT) {
    const double v = partials[i];
    if (MODE == 0) acc = v > acc ? v : acc; else acc += v;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (MODE == 0) acc = warp_max(acc); else acc = warp_sum(acc);
  if (lane == 0) scratch[w] = acc;
  __syncthreads();
  if (w == 0) {
    double r = lane < kHT / 32 ? scratch[lane] : 0.0;
    if (MODE == 0) r = warp_max(r); else r = warp_sum(r);
    if (lane == 0) *out = r;
  }
}

// v[i] /= *norm   (left alone when the norm is zero: an edgeless graph keeps its zeros)
template <typename RT>
__global__ void divide_kernel(RT* v, size_t n, const double* norm)
{
  const RT d = static_cast<RT>(*norm);
  if (!(d > RT(0))) return;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    v[i] = v[i] / d;
}

template <typename RT>
struct vector_ops {
  b200_handle* h;
  unsigned grid;
  dbuf<double> partials;
  dbuf<double> scalar;
  vector_ops(b200_handle* h_, size_t n) : h(h_), grid(hgrid(n, h_->sm_count))
  {
    partials.alloc(grid, h->stream);
    scalar.alloc(1, h->stream);
  }
  template <int MODE>
  void reduce(const RT* a, const RT* b, size_t n)
  {
    reduce_partials_kernel<RT, MODE><<<grid, kHT, 0, h->stream>>>(a, b, n, partials.data());
    fold_partials_kernel<(MODE == 0 ? 0 : 1)><<<1, kHT, 0, h->stream>>>(
      partials.data(), static_cast<int>(grid), scalar.data());
    CUDA_TRY(cudaGetLastError());
  }
  void normalize_max(RT* v, size_t n)
  {
    reduce<0>(v, nullptr, n);
    divide_kernel<RT><<<grid, kHT, 0, h->stream>>>(v, n, scalar.data());
  }
  void normalize_sum(RT* v, size_t n)
  {
    reduce<1>(v, nullptr, n);
    divide_kernel<RT><<<grid, kHT, 0, h->stream>>>(v, n, scalar.data());
  }
  double l1_diff(const RT* a, const RT* b, size_t n)
  {
    reduce<2>(a, b, n);
    double out = 0.0;
    CUDA_TRY(cudaMemcpyAsync(&out, scalar.data(), 8, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return out;
  }
};

// the graph again with rows = `rows_are_dst ? destinations : sources`, internal ids, no weights
std::shared_ptr<b200_graph> make_companion(b200_handle* h, b200_graph* g, bool rows_are_dst)
{
  cudaStream_t s     = h->stream;
  edge_partition& ep = g->parts[0];
  const size_t ne    = static_cast<size_t>(ep.num_edges);
  const size_t nv    = static_cast<size_t>(g->num_vertices);
  dbuf<int32_t> majors(ne, s), minors(ne, s), verts(nv, s);
  if (ne) {
    unsigned grid = static_cast<unsigned>(std::min<size_t>((nv * 32 + kHT - 1) / kHT, 65535u * 8));
    if (grid == 0) grid = 1;
    if (ep.offsets64)
      expand_rows_kernel<int64_t><<<grid, kHT, 0, s>>>(
        reinterpret_cast<const int64_t*>(ep.offsets.data()), ep.indices.data(), ep.num_rows,
        majors.data(), minors.data());
    else
      expand_rows_kernel<int32_t><<<grid, kHT, 0, s>>>(
        reinterpret_cast<const int32_t*>(ep.offsets.data()), ep.indices.data(), ep.num_rows,
        majors.data(), minors.data());
  }
  iota32_kernel<<<hgrid(nv, h->sm_count), kHT, 0, s>>>(verts.data(), nv);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  // g stores rows = destinations: majors are destinations, minors are sources
  const int32_t* src = g->store_transposed ? minors.data() : majors.data();
  const int32_t* dst = g->store_transposed ? majors.data() : minors.data();
  auto c = std::make_shared<b200_graph>();
  graph_create_sg(h, src, dst, nullptr, ne, B200_INT32, B200_FLOAT32, verts.data(), nv,
                  /*store_transposed=*/rows_are_dst, /*renumber=*/false, c.get());
  B200_EXPECTS(c->num_vertices == g->num_vertices, kUnknownError,
               "HITS: companion graph has a different vertex count");
  return c;
}

template <typename RT>
void hits_sg(b200_handle* h, b200_graph* g, const b200_hits_params_t* p, RT* hubs, RT* authorities,
             b200_centrality_stats_t* stats)
{
  cudaStream_t s   = h->stream;
  const size_t n   = static_cast<size_t>(g->num_vertices);
  const int vt     = sizeof(RT) == 4 ? B200_FLOAT32 : B200_FLOAT64;
  const bool fixed = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations : static_cast<int64_t>(p->max_iterations);
  if (stats) *stats = b200_centrality_stats_t{};
  if (n == 0) return;

  if (!g->hits_csr) g->hits_csr = make_companion(h, g, false);
  b200_graph* pull_auth = g;   // rows = destinations
  if (g->weighted) {
    if (!g->hits_csc) g->hits_csc = make_companion(h, g, true);
    pull_auth = g->hits_csc.get();
  }
  b200_graph* pull_hubs = g->hits_csr.get();   // rows = sources

  vector_ops<RT> ops(h, n);
  dbuf<RT> temp(n, s);
  RT* prev = hubs;
  RT* curr = temp.data();
  if (p->has_initial_hubs_guess) {
    ops.normalize_sum(prev, n);   // hits_impl.cuh:91-93
  } else {
    fill_kernel<RT><<<ops.grid, kHT, 0, s>>>(prev, n, RT(1) / static_cast<RT>(g->num_vertices));
  }
  CUDA_TRY(cudaGetLastError());

  cudaEvent_t e0, e1;
  CUDA_TRY(cudaEventCreate(&e0));
  CUDA_TRY(cudaEventCreate(&e1));
  CUDA_TRY(cudaEventRecord(e0, s));
  double diff   = std::numeric_limits<double>::max();
  int64_t final_iteration_count = cap;
  int converged = 0;
  int64_t spmvs = 0;
  // the two launchers (layouts, schedules, scratch) are built once, not per iteration
  struct spmv_guard {
    partition_spmv* p{nullptr};
    ~spmv_guard() { if (p) partition_spmv_free(p); }
  } sa, sh;
  sa.p = partition_spmv_create(h, pull_auth, 0, vt);
  sh.p = partition_spmv_create(h, pull_hubs, 0, vt);
  for (int64_t iter = 0; iter < cap; ++iter) {
    partition_spmv_run(sa.p, prev, authorities, nullptr);   // authorities = A^T hubs
    partition_spmv_run(sh.p, authorities, curr, nullptr);   // hubs = A authorities
    spmvs += 2;
    ops.normalize_max(curr, n);
    ops.normalize_max(authorities, n);
    diff = ops.l1_diff(curr, prev, n);
    std::swap(prev, curr);   // prev holds the latest hubs
    if (!fixed && diff < p->epsilon) {
      final_iteration_count = iter;
      converged             = 1;
      break;
    }
  }
  if (p->normalize) {
    ops.normalize_sum(prev, n);
    ops.normalize_sum(authorities, n);
  }
  if (prev != hubs)
    CUDA_TRY(cudaMemcpyAsync(hubs, prev, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  CUDA_TRY(cudaEventRecord(e1, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  float ms = 0.f;
  CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  if (stats) {
    stats->iterations   = final_iteration_count;
    stats->converged    = converged;
    stats->last_diff    = diff;
    stats->elapsed_ms   = ms;
    stats->gpu_launches = spmvs;   // SpMVs (each 1-3 kernel launches); vertex passes not counted
  }
}

}  // namespace

void hits(b200_handle* h, b200_graph* g, const b200_hits_params_t* p, void* hubs, void* authorities,
          b200_centrality_stats_t* stats)
{
  B200_EXPECTS(g->store_transposed, kInvalidInput,
               "HITS needs a graph stored transposed (pull layout).");
  B200_EXPECTS(!g->multi_gpu, kNotImplemented, "HITS on a multi-GPU graph");
  B200_EXPECTS(p->epsilon >= 0.0, kInvalidInput,
               "Invalid input argument: epsilon should be non-negative.");
  B200_EXPECTS(p->value_type == B200_FLOAT32 || p->value_type == B200_FLOAT64,
               kUnsupportedTypeCombination, "result type must be FLOAT32 or FLOAT64");
  B200_EXPECTS(hubs && authorities, kInvalidInput, "NULL argument");
  if (p->value_type == B200_FLOAT32)
    hits_sg<float>(h, g, p, static_cast<float*>(hubs), static_cast<float*>(authorities), stats);
  else
    hits_sg<double>(h, g, p, static_cast<double*>(hubs), static_cast<double*>(authorities), stats);
}

}  // namespace b200
