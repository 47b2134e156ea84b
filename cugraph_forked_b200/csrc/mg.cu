// Multi-GPU (2D-partitioned) power-iteration drivers.  One iteration, as the reference does it
// (per_v_transform_reduce_incoming_outgoing_e.cuh:831-967, update_edge_src_dst_property.cuh:
// 243-357, pagerank_impl.cuh:209-292):
//   1. allgatherv of the local slice of the gathered vector over row_comm -> minor buffer
//   2. for each of the C local edge partitions: the pull kernel with the raw epilogue, then
//      ncclReduce of its result onto the col_comm rank that owns those majors; the reduction
//      runs on a second stream so it overlaps the next partition's kernel
//   3. one vertex pass over the local vertices (the driver's epilogue) + two scalars
//      allreduced over all ranks; the stop test stays on the device, every rank sees the same
//      scalars and therefore stops at the same iteration
// Every rank calls these functions collectively with the same arguments.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "centrality.cuh"
#include "mg_comm.cuh"
#include "mg_peer.cuh"
#include "prims.cuh"

namespace b200 {

namespace {

constexpr int kVT = 256;

inline unsigned vgrid(size_t n, int sm_count)
{
  size_t b   = (n + kVT * 4 - 1) / (kVT * 4);
  size_t cap = static_cast<size_t>(sm_count) * 8;
  if (b > cap) b = cap;
  if (b == 0) b = 1;
  return static_cast<unsigned>(b);
}

int env_int(const char* name, int dflt)
{
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : dflt;
}

template <typename RT>
struct nccl_type;
template <>
struct nccl_type<float> {
  static constexpr ncclDataType_t value = ncclFloat;
};
template <>
struct nccl_type<double> {
  static constexpr ncclDataType_t value = ncclDouble;
};

// ---- vertex passes ------------------------------------------------------------------------
// partials layout: [2][gridDim.x]
__device__ __forceinline__ void store_partials(double a, double b, double* partials, double* scratch)
{
  a = block_sum(a, scratch);
  b = block_sum(b, scratch);
  if (threadIdx.x == 0) {
    partials[blockIdx.x]             = a;
    partials[gridDim.x + blockIdx.x] = b;
  }
}

template <typename RT, typename WT>
__global__ void __launch_bounds__(kVT) mg_scale_kernel(const WT* outw, size_t n, RT alpha, RT* scale)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    RT w     = static_cast<RT>(outw[i]);
    scale[i] = (w == RT(0)) ? RT(-1) : alpha / w;
  }
}

template <typename RT>
__global__ void __launch_bounds__(kVT) mg_sum_kernel(const RT* v, size_t n, double* partials)
{
  __shared__ double scratch[32];
  double a = 0.0;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    a += static_cast<double>(v[i]);
  store_partials(a, 0.0, partials, scratch);
}

// out2[0] = sum partials[0][*], out2[1] = sum partials[1][*]  (zero once the loop is done)
__global__ void __launch_bounds__(256) mg_fold_kernel(const double* partials, int n,
                                                      const iter_state* st, double* out2)
{
  __shared__ double scratch[32];
  double a = 0.0, b = 0.0;
  if (!(st && st->done)) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      a += partials[i];
      b += partials[n + i];
    }
  }
  a = block_sum(a, scratch);
  b = block_sum(b, scratch);
  if (threadIdx.x == 0) {
    out2[0] = a;
    out2[1] = b;
  }
}

// after the allreduce: publish the scalars and apply the stop test (same on every rank)
__global__ void mg_commit_kernel(const double* in2, iter_state* st, double threshold,
                                 int64_t iteration, int check, int which)
{
  if (st->done) return;
  if (which & 1) st->a = in2[0];
  if (which & 2) st->b = in2[1];
  if (iteration > 0) st->iterations = iteration;
  if (check && in2[1] < threshold) {
    st->converged = 1;
    st->done      = 1;
  }
}

// pr = guess / sum (or 1 / V); x = pr * scale; partial a = dangling mass
template <typename RT>
__global__ void __launch_bounds__(kVT) mg_pagerank_init_kernel(RT* pr, const RT* scale, RT* x,
                                                              size_t n, int has_guess,
                                                              const iter_state* st, RT uniform,
                                                              double* partials)
{
  __shared__ double scratch[32];
  double dang = 0.0;
  RT sum      = has_guess ? static_cast<RT>(st->a) : RT(1);
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    RT v  = has_guess ? pr[i] / sum : uniform;
    pr[i] = v;
    RT sc = scale[i];
    if (sc < RT(0)) {
      dang += static_cast<double>(v);
      x[i] = RT(0);
    } else {
      x[i] = v * sc;
    }
  }
  store_partials(dang, 0.0, partials, scratch);
}

// the PageRank epilogue of epilogues.cuh as a vertex pass (pagerank_impl.cuh:209-292)
template <typename RT>
__global__ void __launch_bounds__(kVT) mg_pagerank_update_kernel(
  RT* pr, const RT* sum, const RT* scale, const RT* pers, RT* x, size_t n, const iter_state* st,
  RT alpha, RT one_minus_alpha, RT num_vertices, double* partials)
{
  __shared__ double scratch[32];
  double dang = 0.0, diff = 0.0;
  if (!st->done) {
    RT dangling  = static_cast<RT>(st->a);
    RT base      = dangling * alpha + one_minus_alpha;
    RT unvarying = pers ? RT(0) : base / num_vertices;
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
      RT v = sum[i] + unvarying;
      if (pers) v += base * pers[i];
      diff += fabs(static_cast<double>(v) - static_cast<double>(pr[i]));
      pr[i] = v;
      RT sc = scale[i];
      if (sc < RT(0)) {
        dang += static_cast<double>(v);
        x[i] = RT(0);
      } else {
        x[i] = v * sc;
      }
    }
  }
  store_partials(dang, diff, partials, scratch);
}

// The same update with the column reduction fused in (peer-memory exchange): the row's sum is
// the sum of the NSRC copies of this rank's partial-sum slice - its own and NSRC - 1 read over
// NVLink (ld.cv: never a stale line) - added in column-rank order; 16-byte loads and stores.
template <typename RT, int NSRC>
__global__ void __launch_bounds__(kVT) mg_pagerank_update_fused_kernel(
  RT* pr, const RT* const* srcs, const RT* scale, const RT* pers, RT* x, RT* x2, size_t n,
  const iter_state* st, RT alpha, RT one_minus_alpha, RT num_vertices, double* partials)
{
  __shared__ double scratch[32];
  constexpr int V = 16 / sizeof(RT);
  double dang = 0.0, diff = 0.0;
  if (!st->done) {
    const RT* s[NSRC];
#pragma unroll
    for (int j = 0; j < NSRC; ++j) s[j] = srcs[j];
    const RT dangling  = static_cast<RT>(st->a);
    const RT base      = dangling * alpha + one_minus_alpha;
    const RT unvarying = pers ? RT(0) : base / num_vertices;
    auto one = [&](RT sum, RT old, RT sc, RT pv, RT& newpr, RT& newx) {
      RT v = sum + unvarying;
      if (pers) v += base * pv;
      diff += fabs(static_cast<double>(v) - static_cast<double>(old));
      newpr = v;
      if (sc < RT(0)) { dang += static_cast<double>(v); newx = RT(0); }
      else newx = v * sc;
    };
    const size_t nv = n / V;
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < nv;
         i += size_t(gridDim.x) * blockDim.x) {
      alignas(16) RT acc[V], old[V], sc[V], pv[V], np[V], nx[V];
#pragma unroll
      for (int k = 0; k < V; ++k) { acc[k] = RT(0); pv[k] = RT(0); }
#pragma unroll
      for (int j = 0; j < NSRC; ++j) {
        const int4 q = __ldcv(reinterpret_cast<const int4*>(s[j]) + i);
        const RT* qv = reinterpret_cast<const RT*>(&q);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] += qv[k];
      }
      *reinterpret_cast<int4*>(old) = __ldcs(reinterpret_cast<const int4*>(pr) + i);
      *reinterpret_cast<int4*>(sc)  = __ldcs(reinterpret_cast<const int4*>(scale) + i);
      if (pers) *reinterpret_cast<int4*>(pv) = __ldcs(reinterpret_cast<const int4*>(pers) + i);
#pragma unroll
      for (int k = 0; k < V; ++k) one(acc[k], old[k], sc[k], pv[k], np[k], nx[k]);
      *(reinterpret_cast<int4*>(pr) + i) = *reinterpret_cast<const int4*>(np);
      *(reinterpret_cast<int4*>(x) + i)  = *reinterpret_cast<const int4*>(nx);
      if (x2) {   // the rank's own piece of the minor-range copy (may be 4/8-byte aligned only)
#pragma unroll
        for (int k = 0; k < V; ++k) x2[i * V + k] = nx[k];
      }
    }
    for (size_t i = nv * V + blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
      RT a = RT(0);
#pragma unroll
      for (int j = 0; j < NSRC; ++j) a += __ldcv(s[j] + i);
      RT np1, nx1;
      one(a, pr[i], scale[i], pers ? pers[i] : RT(0), np1, nx1);
      pr[i] = np1;
      x[i]  = nx1;
      if (x2) x2[i] = nx1;
    }
  }
  store_partials(dang, diff, partials, scratch);
}

// Katz (katz_centrality_impl.cuh:104-148): new = alpha * sum + beta; a = sum of squares
template <typename RT>
__global__ void __launch_bounds__(kVT) mg_katz_update_kernel(const RT* old_c, RT* new_c,
                                                            const RT* sum, const RT* betas,
                                                            size_t n, const iter_state* st,
                                                            RT alpha, RT beta, double* partials)
{
  __shared__ double scratch[32];
  double sq = 0.0, diff = 0.0;
  if (!st->done) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
      RT v     = alpha * sum[i] + (betas ? betas[i] : beta);
      new_c[i] = v;
      diff += fabs(static_cast<double>(v) - static_cast<double>(old_c[i]));
      sq += static_cast<double>(v) * static_cast<double>(v);
    }
  }
  store_partials(sq, diff, partials, scratch);
}

template <typename RT>
__global__ void __launch_bounds__(kVT) mg_sumsq_kernel(const RT* v, size_t n, const iter_state* st,
                                                      double* partials)
{
  __shared__ double scratch[32];
  double sq = 0.0;
  if (!st->done) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x)
      sq += static_cast<double>(v[i]) * static_cast<double>(v[i]);
  }
  store_partials(sq, 0.0, partials, scratch);
}

// eigenvector (eigenvector_centrality_impl.cuh:90-135): new = sum / ||sum||, b = |new - old|
template <typename RT>
__global__ void __launch_bounds__(kVT) mg_eigen_normalize_kernel(const RT* sum, const RT* old_c,
                                                                RT* new_c, size_t n,
                                                                const iter_state* st,
                                                                double* partials)
{
  __shared__ double scratch[32];
  double diff = 0.0;
  if (!st->done) {
    RT norm = static_cast<RT>(sqrt(st->a));
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
         i += size_t(gridDim.x) * blockDim.x) {
      RT v     = sum[i] / norm;
      new_c[i] = v;
      diff += fabs(static_cast<double>(v) - static_cast<double>(old_c[i]));
    }
  }
  store_partials(0.0, diff, partials, scratch);
}

template <typename RT>
__global__ void __launch_bounds__(kVT) mg_fill_kernel(RT* v, size_t n, RT x)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    v[i] = x;
}

template <typename RT>
__global__ void __launch_bounds__(kVT) mg_divide_kernel(RT* v, size_t n, RT d)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    v[i] = v[i] / d;
}

template <typename RT>
__global__ void __launch_bounds__(kVT) mg_scatter_pers_kernel(const int32_t* vertices,
                                                             const RT* values, size_t n,
                                                             RT pers_sum, int64_t first,
                                                             int64_t nloc, RT* dense, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    int64_t v = static_cast<int64_t>(vertices[i]) - first;
    if (v < 0 || v >= nloc) { *bad = 1; continue; }
    atomicAdd(&dense[v], values[i] / pers_sum);
  }
}

// ---- the 2D exchange around the partitions' kernels -------------------------------------------
template <typename RT>
struct mg_engine {
  b200_handle* h;
  b200_graph* g;
  int R, C;
  size_t nloc, n_minor;
  std::vector<partition_spmv*> parts;
  std::vector<dbuf<RT>> partial;      // one result buffer per local edge partition
  dbuf<RT> x_minor;                   // gathered vector over the local minor range
  dbuf<RT> sum;                       // reduced result for the local vertices
  dbuf<RT> sum_padded;                // fused path: this rank's (padded) slice after the reduce-scatter
  // peer-memory exchange (mg_peer.cuh): partial sums and the gathered vector go through an
  // IPC-mapped heap, double-buffered
  bool p2p{false};
  bool fuse_reduce{false};          // the driver's update kernel sums the C copies itself
  int last_buf{0};                  // the heap buffer the last SpMV wrote
  peer_heap* heap{nullptr};
  size_t off_part[2]{0, 0}, off_x[2]{0, 0};
  int buf{0};
  dbuf<uint64_t*> d_slots_row, d_slots_col;
  dbuf<const RT*> d_srcs[2];
  dbuf<double*> d_val_slots[2];
  dbuf<uint64_t*> d_flag_slots;
  RT* part_local(int b) const { return heap->at<RT>(h->rank, off_part[b]); }
  // where the next gathered vector goes when the driver publishes it itself: the heap buffer the
  // next spmv() exchanges, and this rank's piece of the minor-range copy
  RT* x_pub() const { return heap->at<RT>(h->rank, off_x[buf]); }
  RT* x_own()
  {
    size_t d = 0;
    for (int j = 0; j < h->row_rank; ++j) d += row_counts[j];
    return x_minor.data() + d;
  }
  void peer_barrier(bool col)
  {
    const uint64_t epoch = col ? ++heap->epoch_col : ++heap->epoch_row;
    peer_barrier_kernel<<<1, 32, 0, h->stream>>>(
      col ? d_slots_col.data() : d_slots_row.data(),
      heap->at<uint64_t>(h->rank, peer_heap::flag_offset(col, 0)), col ? C : R, epoch, heap->d_error);
    CUDA_TRY(cudaGetLastError());
    ++launches;
  }
  void setup_p2p()
  {
    cudaStream_t s = h->stream;
    // one layout for every rank of the box: sizes are the maxima over the world
    std::vector<int64_t> all = allgather_host(
      h, h->world, h->size, {static_cast<int64_t>(g->fused_slice) * C, static_cast<int64_t>(nloc)});
    int64_t max_part = 0, max_x = 0;
    for (int p = 0; p < h->size; ++p) {
      max_part = std::max(max_part, all[2 * p]);
      max_x    = std::max(max_x, all[2 * p + 1]);
    }
    const size_t pb = (static_cast<size_t>(max_part) * sizeof(RT) + 255) & ~size_t{255};
    const size_t xb = (static_cast<size_t>(max_x) * sizeof(RT) + 255) & ~size_t{255};
    heap        = peer_heap_get(h, 2 * pb + 2 * xb);
    if (!heap) {   // agreed by all ranks: fall back to the collective exchange
      p2p = false;
      return;
    }
    off_part[0] = kPeerHeader;
    off_part[1] = kPeerHeader + pb;
    off_x[0]    = kPeerHeader + 2 * pb;
    off_x[1]    = kPeerHeader + 2 * pb + xb;
    // world rank of row peer j / column peer t (partition_manager.hpp:86-103)
    auto row_peer = [&](int j) { return (h->rank / R) * R + j; };
    auto col_peer = [&](int t) { return t * R + (h->rank % R); };
    std::vector<uint64_t*> sr(R), sc(C);
    for (int j = 0; j < R; ++j)
      sr[j] = heap->at<uint64_t>(row_peer(j), peer_heap::flag_offset(false, h->row_rank));
    for (int t = 0; t < C; ++t)
      sc[t] = heap->at<uint64_t>(col_peer(t), peer_heap::flag_offset(true, h->col_rank));
    d_slots_row.alloc(R, s);
    d_slots_col.alloc(C, s);
    CUDA_TRY(cudaMemcpyAsync(d_slots_row.data(), sr.data(), R * sizeof(void*), cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_slots_col.data(), sc.data(), C * sizeof(void*), cudaMemcpyHostToDevice, s));
    for (int b = 0; b < 2; ++b) {
      std::vector<const RT*> src(C);
      for (int t = 0; t < C; ++t)
        src[t] = heap->at<RT>(col_peer(t), off_part[b]) + static_cast<size_t>(h->col_rank) * g->fused_slice;
      d_srcs[b].alloc(C, s);
      CUDA_TRY(cudaMemcpyAsync(d_srcs[b].data(), src.data(), C * sizeof(void*), cudaMemcpyHostToDevice, s));
      CUDA_TRY(cudaStreamSynchronize(s));
    }
    {
      std::vector<uint64_t*> fs(h->size);
      for (int t = 0; t < h->size; ++t) fs[t] = heap->at<uint64_t>(t, peer_heap::scal_flag_offset(h->rank));
      d_flag_slots.alloc(h->size, s);
      CUDA_TRY(cudaMemcpyAsync(d_flag_slots.data(), fs.data(), h->size * sizeof(void*), cudaMemcpyHostToDevice, s));
      for (int par = 0; par < 2; ++par) {
        std::vector<double*> vs(h->size);
        for (int t = 0; t < h->size; ++t) vs[t] = heap->at<double>(t, peer_heap::scal_val_offset(par, h->rank));
        d_val_slots[par].alloc(h->size, s);
        CUDA_TRY(cudaMemcpyAsync(d_val_slots[par].data(), vs.data(), h->size * sizeof(void*),
                                 cudaMemcpyHostToDevice, s));
        CUDA_TRY(cudaStreamSynchronize(s));
      }
    }
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  void check_peer_error()
  {
    if (!p2p) return;
    int e = 0;
    CUDA_TRY(cudaMemcpyAsync(&e, heap->d_error, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    B200_EXPECTS(e == 0, kUnknownError, "multi-GPU exchange: a peer did not reach the barrier (timeout)");
  }
  std::vector<size_t> row_counts;     // vertex partition sizes of the row_comm peers
  dbuf<iter_state> state;
  dbuf<double> partials;              // [2][vg]
  dbuf<double> scal;                  // 2 scalars in flight through the allreduce
  unsigned vg;
  std::vector<cudaEvent_t> ev_part;
  cudaEvent_t ev_red{};
  cudaEvent_t ev_iter{}, ev_copied{};
  int64_t launches{0};
  bool fused{false};
  // device-time of the partitions' kernels on the main stream
  std::vector<cudaEvent_t> ev_time;
  int timed{0};
  // B200_MG_TRACE=1: per-phase device times of the first iterations, printed by rank 0
  bool trace{false};
  std::vector<cudaEvent_t> ev_trace;  // 6 per iteration: start, gathered, kernels, reduced, updated, end
  void mark(int k)
  {
    if (!trace || ev_trace.size() >= 6 * 64 + 6) return;
    if (k == 0 && ev_trace.size() % 6 != 0) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, h->stream);
    ev_trace.push_back(e);
  }
  void print_trace()
  {
    if (!trace || (h->rank != 0 && env_int("B200_MG_TRACE", 0) < 2)) return;
    size_t iters = ev_trace.size() / 6;
    if (iters == 0) return;
    double acc[5] = {0, 0, 0, 0, 0};
    for (size_t i = 0; i < iters; ++i)
      for (int k = 0; k < 5; ++k) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev_trace[6 * i + k], ev_trace[6 * i + k + 1]);
        acc[k] += ms;
      }
    fprintf(stderr,
            "[b200 mg trace] rank %d: %zu iterations: allgather %.1f us, partition kernels %.1f us, "
            "reduce tail %.1f us, vertex pass %.1f us, scalars %.1f us\n",
            h->rank, iters, 1e3 * acc[0] / iters, 1e3 * acc[1] / iters, 1e3 * acc[2] / iters,
            1e3 * acc[3] / iters, 1e3 * acc[4] / iters);
  }

  mg_engine(b200_handle* h_, b200_graph* g_) : h(h_), g(g_)
  {
    cudaStream_t s = h->stream;
    R              = h->row_size;
    C              = h->col_size;
    nloc           = static_cast<size_t>(g->local_last - g->local_first);
    n_minor        = static_cast<size_t>(g->parts[0].num_minors);
    int vt         = sizeof(RT) == 4 ? B200_FLOAT32 : B200_FLOAT64;
    // One launch train over all local partitions (B200_MG_FUSED, also read at graph creation).
    // Round 1 measured it against one launch per partition whose reduction overlaps the next
    // kernel and kept it opt-in (its single-phase kernel was 17 % faster at 8 GPUs / scale 24,
    // but the C reductions no longer overlapped: 545 vs 585 GTEPS).  With the round-2 layout
    // (sell.cuh) the per-partition rectangles of an 8-GPU grid are too small to fill the
    // kernels (3 launches x C per iteration at the efficiency of a scale-20 graph), so the
    // fused rectangle is the default; profiles/r2_notes.md has the per-phase times.
    fused = !g->fused.empty() && env_int("B200_MG_FUSED", 1) != 0;
    if (fused) {
      // one launch over all local partitions; partial sums of partition i are the rows
      // [fused_row_first[i], fused_row_first[i + 1]) of one result buffer
      parts.assign(1, partition_spmv_create(h, g, -1, vt));
      partial.resize(1);
      sum_padded.alloc(static_cast<size_t>(g->fused_slice), s, 8);
      // the peer reduction is instantiated for 2, 4 and 8 column peers (the grids of 2 / 4 / 8 /
      // 16 GPUs); other grids keep the NCCL exchange
      p2p = env_int("B200_MG_P2P", 1) != 0 && (C == 2 || C == 4 || C == 8);
      if (p2p) setup_p2p();
      if (!p2p) partial[0].alloc(static_cast<size_t>(g->fused[0].num_rows), s, 8);
    } else {
      parts.assign(C, nullptr);
      for (int i = 0; i < C; ++i) parts[i] = partition_spmv_create(h, g, i, vt);
      partial.resize(C);
      for (int i = 0; i < C; ++i) partial[i].alloc(static_cast<size_t>(g->parts[i].num_rows), s, 8);
    }
    x_minor.alloc(n_minor, s, 8);
    sum.alloc(nloc, s, 8);
    auto const& vpo = g->vertex_partition_offsets;
    row_counts.resize(R);
    for (int j = 0; j < R; ++j)
      row_counts[j] = static_cast<size_t>(vpo[h->col_rank * R + j + 1] - vpo[h->col_rank * R + j]);
    state.alloc(1, s);
    CUDA_TRY(cudaMemsetAsync(state.data(), 0, sizeof(iter_state), s));
    vg = vgrid(nloc, h->sm_count);
    partials.alloc(2 * vg, s);
    scal.alloc(2, s);
    ev_part.resize(C);
    for (auto& e : ev_part) CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&ev_red, cudaEventDisableTiming));
    trace = env_int("B200_MG_TRACE", 0) != 0;
  }
  ~mg_engine()
  {
    for (auto e : ev_trace) cudaEventDestroy(e);
    for (auto p : parts) partition_spmv_free(p);
    for (auto e : ev_part) cudaEventDestroy(e);
    for (auto e : ev_time) cudaEventDestroy(e);
    if (ev_red) cudaEventDestroy(ev_red);
    if (ev_iter) cudaEventDestroy(ev_iter);
    if (ev_copied) cudaEventDestroy(ev_copied);
    if (ev_fold) cudaEventDestroy(ev_fold);
    if (ev_commit) cudaEventDestroy(ev_commit);
  }

  // sum[local v] = sum over all edges into v of x[src], x given for the local vertices
  void spmv(const RT* x_local)
  {
    cudaStream_t s = h->stream;
    mark(0);
    if (p2p && R > 1) {
      // publish the local slice, wait for the row peers, copy their slices out of their heaps
      std::vector<size_t> dis(R + 1, 0);
      for (int j = 0; j < R; ++j) dis[j + 1] = dis[j] + row_counts[j];
      RT* mine = heap->at<RT>(h->rank, off_x[buf]);
      // published: the driver's update kernel wrote this slice into the heap (and into the local
      // minor-range copy) itself, and the scalar exchange that followed it was a barrier over
      // all ranks - nothing to copy locally, nothing to wait for
      const bool published = x_local == mine;
      if (!published) {
        if (nloc) CUDA_TRY(cudaMemcpyAsync(mine, x_local, nloc * sizeof(RT), cudaMemcpyDeviceToDevice, s));
        peer_barrier(false);
      }
      // published: the peers' slices are copied on the second stream while phase 1 already runs
      // the bands of this rank's own slice (the launcher splits its unit list at the slice)
      const bool overlap = published && env_int("B200_MG_OVERLAP_GATHER", 1) != 0 && h->comm_stream;
      cudaStream_t cs    = overlap ? h->comm_stream : s;
      if (overlap) {
        if (!ev_iter) {
          CUDA_TRY(cudaEventCreateWithFlags(&ev_iter, cudaEventDisableTiming));
          CUDA_TRY(cudaEventCreateWithFlags(&ev_copied, cudaEventDisableTiming));
        }
        CUDA_TRY(cudaEventRecord(ev_iter, s));          // behind the scalar exchange = the barrier
        CUDA_TRY(cudaStreamWaitEvent(cs, ev_iter, 0));
      }
      for (int j = 0; j < R; ++j) {
        if (!row_counts[j]) continue;
        if (j == h->row_rank && published) continue;
        const RT* from = j == h->row_rank ? x_local : heap->at<RT>((h->rank / R) * R + j, off_x[buf]);
        CUDA_TRY(cudaMemcpyAsync(x_minor.data() + dis[j], from, row_counts[j] * sizeof(RT),
                                 cudaMemcpyDefault, cs));
      }
      if (overlap) {
        CUDA_TRY(cudaEventRecord(ev_copied, cs));
        partition_spmv_split(parts[0], static_cast<int64_t>(dis[h->row_rank]),
                             static_cast<int64_t>(dis[h->row_rank + 1]), ev_copied);
      }
    } else if (R > 1) {
      allgatherv(s, h->row, R, h->row_rank, x_local, x_minor.data(), row_counts, sizeof(RT));
    }
    // a 1 x C grid: the minor range IS the local vertex range, the kernels read x_local itself
    const RT* xm = R == 1 ? x_local : x_minor.data();
    mark(1);
    const bool time_it = timed < 256;
    if (time_it) {
      cudaEvent_t a, b;
      CUDA_TRY(cudaEventCreate(&a));
      CUDA_TRY(cudaEventCreate(&b));
      ev_time.push_back(a);
      ev_time.push_back(b);
      CUDA_TRY(cudaEventRecord(a, s));
    }
    if (fused) {
      RT* part = p2p ? part_local(buf) : partial[0].data();
      partition_spmv_run(parts[0], xm, part, state.data());
      if (time_it) {
        CUDA_TRY(cudaEventRecord(ev_time[2 * timed + 1], s));
        ++timed;
      }
      mark(2);
      if (p2p) {
        // every column peer has written its partial sums: sum the C copies of this rank's slice
        // (C - 1 of them over NVLink) in column-rank order
        peer_barrier(true);
        last_buf = buf;
        if (fuse_reduce) {
          buf ^= 1;
          mark(3);
          return;
        }
        const size_t n     = static_cast<size_t>(g->fused_slice);
        const unsigned grd = static_cast<unsigned>(std::min<size_t>((n / (16 / sizeof(RT)) + 255) / 256 + 1,
                                                                    size_t(h->sm_count) * 8));
        if (C == 2)      peer_reduce_kernel<RT, 2><<<grd, 256, 0, s>>>(d_srcs[buf].data(), sum_padded.data(), n);
        else if (C == 4) peer_reduce_kernel<RT, 4><<<grd, 256, 0, s>>>(d_srcs[buf].data(), sum_padded.data(), n);
        else if (C == 8) peer_reduce_kernel<RT, 8><<<grd, 256, 0, s>>>(d_srcs[buf].data(), sum_padded.data(), n);
        else B200_EXPECTS(false, kNotImplemented, "peer reduction: unsupported column count");
        CUDA_TRY(cudaGetLastError());
        ++launches;
        buf ^= 1;
        mark(3);
        return;
      }
      // B200_MG_P2P=0: ONE reduce-scatter on the column communicator (the C slices of the fused
      // result have one padded size)
      NCCL_TRY(ncclReduceScatter(partial[0].data(), sum_padded.data(),
                                 static_cast<size_t>(g->fused_slice), nccl_type<RT>::value, ncclSum,
                                 h->col, s));
      mark(3);
      return;
    }
    for (int i = 0; i < C; ++i) {
      partition_spmv_run(parts[i], xm, partial[i].data(), state.data());
      if (C == 1) break;
      CUDA_TRY(cudaEventRecord(ev_part[i], s));
      CUDA_TRY(cudaStreamWaitEvent(h->comm_stream, ev_part[i], 0));
      size_t cnt = static_cast<size_t>(g->parts[i].num_rows);
      if (cnt)
        NCCL_TRY(ncclReduce(partial[i].data(), sum.data(), cnt, nccl_type<RT>::value, ncclSum, i,
                            h->col, h->comm_stream));
    }
    if (time_it) {
      CUDA_TRY(cudaEventRecord(ev_time[2 * timed + 1], s));
      ++timed;
    }
    mark(2);
    if (C > 1) {
      CUDA_TRY(cudaEventRecord(ev_red, h->comm_stream));
      CUDA_TRY(cudaStreamWaitEvent(s, ev_red, 0));
    }
    mark(3);
  }
  const RT* result() const { return C == 1 ? partial[0].data() : (fused ? sum_padded.data() : sum.data()); }

  // scal = allreduce(fold(partials)); then commit into the state.  With `deferred` the
  // allreduce and the commit run on the second stream and the caller joins them with
  // join_scalars() right before the next consumer of the state (the next vertex pass): the
  // ~100 us small-message latency then hides behind the next iteration's exchange and kernels.
  cudaEvent_t ev_fold{}, ev_commit{};
  bool pending_join{false};
  void reduce_scalars(double threshold, int64_t iteration, bool check, int which,
                      bool deferred = false)
  {
    cudaStream_t s = h->stream;
    if (iteration > 0) mark(4);
    if (p2p) {
      const uint64_t epoch = ++heap->epoch_scal;
      const int par        = static_cast<int>(epoch & 1);
      peer_scalars_kernel<<<1, 256, 0, s>>>(
        partials.data(), static_cast<int>(vg), d_val_slots[par].data(), d_flag_slots.data(),
        heap->at<double>(h->rank, peer_heap::scal_val_offset(par, 0)),
        heap->at<uint64_t>(h->rank, peer_heap::scal_flag_offset(0)), h->size, epoch, state.data(),
        threshold, iteration, check ? 1 : 0, which, heap->d_error);
      CUDA_TRY(cudaGetLastError());
      launches += 1;
      if (iteration > 0) mark(5);
      return;
    }
    mg_fold_kernel<<<1, 256, 0, s>>>(partials.data(), static_cast<int>(vg), state.data(), scal.data());
    CUDA_TRY(cudaGetLastError());
    cudaStream_t cs = s;
    if (deferred && h->size > 1) {
      if (!ev_fold) {
        CUDA_TRY(cudaEventCreateWithFlags(&ev_fold, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&ev_commit, cudaEventDisableTiming));
      }
      cs = h->comm_stream;
      CUDA_TRY(cudaEventRecord(ev_fold, s));
      CUDA_TRY(cudaStreamWaitEvent(cs, ev_fold, 0));
    }
    if (h->size > 1)
      NCCL_TRY(ncclAllReduce(scal.data(), scal.data(), 2, ncclDouble, ncclSum, h->world, cs));
    mg_commit_kernel<<<1, 1, 0, cs>>>(scal.data(), state.data(), threshold, iteration, check ? 1 : 0,
                                      which);
    CUDA_TRY(cudaGetLastError());
    if (cs != s) {
      CUDA_TRY(cudaEventRecord(ev_commit, cs));
      pending_join = true;
    }
    launches += 2;
    if (iteration > 0) mark(5);
  }
  void join_scalars()
  {
    if (!pending_join) return;
    CUDA_TRY(cudaStreamWaitEvent(h->stream, ev_commit, 0));
    pending_join = false;
  }

  iter_state read_state()
  {
    join_scalars();
    iter_state hs;
    CUDA_TRY(cudaMemcpyAsync(&hs, state.data(), sizeof hs, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    check_peer_error();
    return hs;
  }

  void fill_stats(b200_centrality_stats_t* st, iter_state const& hs, double ms)
  {
    print_trace();
    if (!st) return;
    st->iterations   = hs.iterations;
    st->converged    = hs.converged;
    st->last_diff    = hs.b;
    st->elapsed_ms   = ms;
    int64_t l        = launches;
    for (auto p : parts) l += partition_spmv_launches(p);
    st->gpu_launches = l;
    double tot       = 0.0;
    for (int i = 0; i < timed; ++i) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev_time[2 * i], ev_time[2 * i + 1]);
      tot += t;
    }
    st->spmv_ms       = tot;
    st->spmv_launches = timed;
  }
};

struct scoped_timer {
  cudaEvent_t e0{}, e1{};
  cudaStream_t s;
  explicit scoped_timer(cudaStream_t s_) : s(s_)
  {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, s);
  }
  ~scoped_timer()
  {
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  double stop()
  {
    cudaEventRecord(e1, s);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  }
};

}  // namespace

// ------------------------------------------------------------------------------------------
// PageRank
// ------------------------------------------------------------------------------------------
template <typename RT>
static void pagerank_mg_t(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, RT* pr_user,
                          b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const RT alpha    = static_cast<RT>(p->alpha);
  const RT epsilon  = static_cast<RT>(p->epsilon);
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  if (stats) *stats = b200_centrality_stats_t{};
  if (g->num_vertices == 0) return;

  mg_engine<RT> E(h, g);
  const size_t n = E.nloc;
  dbuf<RT> scale(n, s, 8), pr(n, s, 8), x(n, s, 8);
  if (p->has_initial_guess && n)
    CUDA_TRY(cudaMemcpyAsync(pr.data(), pr_user, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  if (p->precomputed_out_weight_sums) {
    mg_scale_kernel<RT, RT><<<E.vg, kVT, 0, s>>>(
      static_cast<const RT*>(p->precomputed_out_weight_sums), n, alpha, scale.data());
  } else {
    graph_out_weight_sums(h, g);
    mg_scale_kernel<RT, double><<<E.vg, kVT, 0, s>>>(g->out_weight_sums.data(), n, alpha, scale.data());
  }
  CUDA_TRY(cudaGetLastError());

  // personalization: the caller's pairs are for vertices of this rank's partition (global
  // internal ids); the normalising sum is global (pagerank_impl.cuh:111-150)
  dbuf<RT> pers;
  const bool has_pers_arg = p->personalization_vertices != nullptr;
  if (has_pers_arg)
    B200_EXPECTS(p->personalization_values != nullptr, kInvalidInput,
                 "Invalid input argument: if personalization_vertices.has_value() is true, "
                 "personalization_values.has_value() and personalization_vector_size.has_value() "
                 "should be true as well.");
  // every rank takes the same branch: the vector is in use when any rank passes entries
  int64_t np_global = 0;
  {
    std::vector<int64_t> all =
      allgather_host(h, h->world, h->size, {static_cast<int64_t>(p->personalization_size)});
    for (int64_t v : all) np_global += v;
  }
  const bool personalized = np_global > 0;
  if (personalized) {
    size_t np = static_cast<size_t>(p->personalization_size);
    unsigned pg = vgrid(np, h->sm_count);
    dbuf<double> pparts(2 * std::max(pg, E.vg), s);
    mg_sum_kernel<RT><<<pg, kVT, 0, s>>>(static_cast<const RT*>(p->personalization_values), np,
                                         pparts.data());
    mg_fold_kernel<<<1, 256, 0, s>>>(pparts.data(), static_cast<int>(pg), nullptr, E.scal.data());
    CUDA_TRY(cudaGetLastError());
    if (h->size > 1)
      NCCL_TRY(ncclAllReduce(E.scal.data(), E.scal.data(), 2, ncclDouble, ncclSum, h->world, s));
    double h2[2];
    CUDA_TRY(cudaMemcpyAsync(h2, E.scal.data(), 16, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    RT pers_sum = static_cast<RT>(h2[0]);
    B200_EXPECTS(pers_sum > RT(0), kInvalidInput,
                 "Invalid input argument: sum of personalization valuese should be positive.");
    pers.alloc(n, s);
    CUDA_TRY(cudaMemsetAsync(pers.data(), 0, std::max<size_t>(n, 1) * sizeof(RT), s));
    dbuf<int> bad(1, s);
    CUDA_TRY(cudaMemsetAsync(bad.data(), 0, 4, s));
    if (np) {
      mg_scatter_pers_kernel<RT><<<pg, kVT, 0, s>>>(
        static_cast<const int32_t*>(p->personalization_vertices),
        static_cast<const RT*>(p->personalization_values), np, pers_sum, g->local_first,
        static_cast<int64_t>(n), pers.data(), bad.data());
      CUDA_TRY(cudaGetLastError());
    }
    int h_bad = 0;
    CUDA_TRY(cudaMemcpyAsync(&h_bad, bad.data(), 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    std::vector<int64_t> flags = allgather_host(h, h->world, h->size, {static_cast<int64_t>(h_bad)});
    for (int64_t f : flags)
      B200_EXPECTS(!f, kInvalidInput,
                   "Invalid input argument: peresonalization vertices have invalid vertex IDs.");
  }

  // initial values
  if (p->has_initial_guess) {
    mg_sum_kernel<RT><<<E.vg, kVT, 0, s>>>(pr.data(), n, E.partials.data());
    CUDA_TRY(cudaGetLastError());
    E.reduce_scalars(0.0, 0, false, 1);
    iter_state hs = E.read_state();
    B200_EXPECTS(static_cast<RT>(hs.a) > RT(0), kInvalidInput,
                 "Invalid input argument: sum of the PageRank initial guess values should be "
                 "positive.");
  }
  mg_pagerank_init_kernel<RT><<<E.vg, kVT, 0, s>>>(
    pr.data(), scale.data(), x.data(), n, p->has_initial_guess ? 1 : 0, E.state.data(),
    RT(1) / static_cast<RT>(g->num_vertices), E.partials.data());
  CUDA_TRY(cudaGetLastError());
  E.reduce_scalars(0.0, 0, false, 1);

  // peer-memory exchange: the update kernel adds the C copies of the partial sums itself
  E.fuse_reduce = E.p2p && (E.C == 2 || E.C == 4 || E.C == 8) && env_int("B200_MG_FUSE_UPDATE", 1) != 0;
  const int check_every = std::max(1, env_int("B200_CHECK_EVERY", 8));
  scoped_timer timer(s);
  const auto host_t0 = std::chrono::steady_clock::now();
  iter_state hs{};
  const bool publish = E.fuse_reduce && E.R > 1 && env_int("B200_MG_PUBLISH_X", 1) != 0;
  const RT* x_cur    = x.data();
  for (int64_t it = 0; it < cap; ++it) {
    E.spmv(x_cur);
    E.join_scalars();  // the dangling mass and the stop flag of the previous iteration
    RT* x_out = publish ? E.x_pub() : x.data();
    RT* x_two = publish ? E.x_own() : nullptr;
    x_cur     = x_out;
    if (E.fuse_reduce) {
      auto go = [&](auto tag) {
        constexpr int NS = decltype(tag)::value;
        mg_pagerank_update_fused_kernel<RT, NS><<<E.vg, kVT, 0, s>>>(
          pr.data(), E.d_srcs[E.last_buf].data(), scale.data(), personalized ? pers.data() : nullptr,
          x_out, x_two, n, E.state.data(), alpha, static_cast<RT>(1.0 - p->alpha),
          static_cast<RT>(g->num_vertices), E.partials.data());
      };
      if (E.C == 2) go(std::integral_constant<int, 2>{});
      else if (E.C == 4) go(std::integral_constant<int, 4>{});
      else go(std::integral_constant<int, 8>{});
    } else {
      mg_pagerank_update_kernel<RT><<<E.vg, kVT, 0, s>>>(
        pr.data(), E.result(), scale.data(), personalized ? pers.data() : nullptr, x.data(), n,
        E.state.data(), alpha, static_cast<RT>(1.0 - p->alpha), static_cast<RT>(g->num_vertices),
        E.partials.data());
    }
    CUDA_TRY(cudaGetLastError());
    ++E.launches;
    E.reduce_scalars(static_cast<double>(epsilon), it + 1, !fixed, 3, true);
    if (!fixed && ((it + 1) % check_every == 0)) {
      hs = E.read_state();
      if (hs.done) break;
    }
  }
  const double host_ms =
    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
  E.join_scalars();
  double ms = timer.stop();
  if (E.trace && h->rank == 0)
    fprintf(stderr, "[b200 mg trace] host enqueue %.3f ms, device %.3f ms for the loop\n", host_ms, ms);
  hs        = E.read_state();
  E.fill_stats(stats, hs, ms);
  if (n) CUDA_TRY(cudaMemcpyAsync(pr_user, pr.data(), n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  if (!fixed && !hs.converged) throw error(kUnknownError, "PageRank failed to converge.");
}

void pagerank_mg(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, void* pageranks,
                 b200_centrality_stats_t* stats)
{
  B200_EXPECTS(h->world != nullptr, kInvalidHandle, "multi-GPU PageRank needs an MG handle");
  if (p->value_type == B200_FLOAT32)
    pagerank_mg_t<float>(h, g, p, static_cast<float*>(pageranks), stats);
  else
    pagerank_mg_t<double>(h, g, p, static_cast<double*>(pageranks), stats);
}

// ------------------------------------------------------------------------------------------
// Katz centrality
// ------------------------------------------------------------------------------------------
template <typename RT>
static void katz_mg_t(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, RT* kc,
                      b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  if (stats) *stats = b200_centrality_stats_t{};
  if (g->num_vertices == 0) return;
  mg_engine<RT> E(h, g);
  const size_t n = E.nloc;
  dbuf<RT> work0(n, s, 8), work1(n, s, 8);
  if (p->has_initial_guess && n)
    CUDA_TRY(cudaMemcpyAsync(work0.data(), kc, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  else
    CUDA_TRY(cudaMemsetAsync(work0.data(), 0, std::max<size_t>(n, 1) * sizeof(RT), s));
  RT* buf[2] = {work0.data(), work1.data()};
  const int check_every = std::max(1, env_int("B200_CHECK_EVERY", 8));
  scoped_timer timer(s);
  iter_state hs{};
  for (int64_t it = 0; it < cap; ++it) {
    RT* old_c = buf[it & 1];
    RT* new_c = buf[(it + 1) & 1];
    E.spmv(old_c);
    mg_katz_update_kernel<RT><<<E.vg, kVT, 0, s>>>(
      old_c, new_c, E.result(), static_cast<const RT*>(p->betas), n, E.state.data(),
      static_cast<RT>(p->alpha), static_cast<RT>(p->beta), E.partials.data());
    CUDA_TRY(cudaGetLastError());
    ++E.launches;
    E.reduce_scalars(static_cast<double>(static_cast<RT>(p->epsilon)), it + 1, !fixed, 3);
    if (!fixed && ((it + 1) % check_every == 0)) {
      hs = E.read_state();
      if (hs.done) break;
    }
  }
  double ms = timer.stop();
  hs        = E.read_state();
  E.fill_stats(stats, hs, ms);
  if (!fixed && !hs.converged) throw error(kUnknownError, "Katz Centrality failed to converge.");
  RT* result = buf[hs.iterations & 1];
  if (n) CUDA_TRY(cudaMemcpyAsync(kc, result, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  if (p->normalize) {
    RT l2 = static_cast<RT>(std::sqrt(static_cast<RT>(hs.a)));
    B200_EXPECTS(l2 > RT(0), kUnknownError,
                 "L2 norm of the computed Katz Centrality values should be positive.");
    if (n) {
      mg_divide_kernel<RT><<<E.vg, kVT, 0, s>>>(kc, n, l2);
      CUDA_TRY(cudaGetLastError());
    }
  }
  CUDA_TRY(cudaStreamSynchronize(s));
}

void katz_mg(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, void* out,
             b200_centrality_stats_t* stats)
{
  B200_EXPECTS(h->world != nullptr, kInvalidHandle, "multi-GPU Katz needs an MG handle");
  if (p->value_type == B200_FLOAT32) katz_mg_t<float>(h, g, p, static_cast<float*>(out), stats);
  else katz_mg_t<double>(h, g, p, static_cast<double*>(out), stats);
}

// ------------------------------------------------------------------------------------------
// eigenvector centrality
// ------------------------------------------------------------------------------------------
template <typename RT>
static void eigenvector_mg_t(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p,
                             RT* c, b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  if (stats) *stats = b200_centrality_stats_t{};
  if (g->num_vertices == 0) return;
  mg_engine<RT> E(h, g);
  const size_t n = E.nloc;
  dbuf<RT> work0(n, s, 8), work1(n, s, 8);
  if (p->has_initial_guess && n) {
    CUDA_TRY(cudaMemcpyAsync(work0.data(), c, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  } else {
    mg_fill_kernel<RT><<<E.vg, kVT, 0, s>>>(work0.data(), n, RT(1) / static_cast<RT>(g->num_vertices));
    CUDA_TRY(cudaGetLastError());
  }
  RT* buf[2] = {work0.data(), work1.data()};
  const double threshold =
    static_cast<double>(static_cast<RT>(g->num_vertices) * static_cast<RT>(p->epsilon));
  const int check_every = std::max(1, env_int("B200_CHECK_EVERY", 8));
  scoped_timer timer(s);
  iter_state hs{};
  for (int64_t it = 0; it < cap; ++it) {
    RT* old_c = buf[it & 1];
    RT* new_c = buf[(it + 1) & 1];
    E.spmv(old_c);
    mg_sumsq_kernel<RT><<<E.vg, kVT, 0, s>>>(E.result(), n, E.state.data(), E.partials.data());
    CUDA_TRY(cudaGetLastError());
    E.reduce_scalars(0.0, 0, false, 1);
    mg_eigen_normalize_kernel<RT><<<E.vg, kVT, 0, s>>>(E.result(), old_c, new_c, n, E.state.data(),
                                                       E.partials.data());
    CUDA_TRY(cudaGetLastError());
    E.launches += 2;
    E.reduce_scalars(threshold, it + 1, !fixed, 2);
    if (!fixed && ((it + 1) % check_every == 0)) {
      hs = E.read_state();
      if (hs.done) break;
    }
  }
  double ms = timer.stop();
  hs        = E.read_state();
  E.fill_stats(stats, hs, ms);
  if (!fixed && !hs.converged)
    throw error(kUnknownError, "Eigenvector Centrality failed to converge.");
  RT* result = buf[hs.iterations & 1];
  if (n) CUDA_TRY(cudaMemcpyAsync(c, result, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
}

void eigenvector_mg(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p, void* out,
                    b200_centrality_stats_t* stats)
{
  B200_EXPECTS(h->world != nullptr, kInvalidHandle, "multi-GPU eigenvector needs an MG handle");
  if (p->value_type == B200_FLOAT32)
    eigenvector_mg_t<float>(h, g, p, static_cast<float*>(out), stats);
  else
    eigenvector_mg_t<double>(h, g, p, static_cast<double*>(out), stats);
}

}  // namespace b200
