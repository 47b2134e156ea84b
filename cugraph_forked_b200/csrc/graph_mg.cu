// Multi-GPU (2D-partitioned) graph build.  B200-native replacement of
//   cpp/src/detail/shuffle_wrappers.cu:41-174 (edges to their 2D owner)
//   cpp/src/detail/graph_utils.cuh:45-138 (MurmurHash3 vertex / edge -> rank functors)
//   cpp/src/structure/renumber_edgelist_impl.cuh:96-452, :665-921 (MG renumbering)
//   cpp/src/structure/create_graph_from_edgelist_impl.cuh:207-555 (per-partition compression)
//   cpp/include/cugraph/graph_view.hpp:81-242 (partition_t ranges)
// One process per GPU; all exchanges are NCCL send/recv or broadcast groups on the handle's
// stream (mg_comm.cuh).  Where the reference builds cuco hash maps to translate ids, this
// build sorts the gathered label arrays once and binary-searches them; where it compresses
// each partition with atomics and a segmented sort, one stable radix sort of
// (concatenated major row << 32 | minor offset) keys lays out all local partitions at once.
#include <cstdlib>

#include "graph_kernels.cuh"
#include "mg_comm.cuh"

namespace b200 {

namespace {

// MurmurHash3_x86_32, seed 0, over the 4 or 8 bytes of the key: the published algorithm that
// cuco::detail::MurmurHash3_32<vertex_t> implements (graph_utils.cuh:50; parity unpinned, see
// oracle/graph_oracle.py)
__device__ __forceinline__ uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
__device__ __forceinline__ uint32_t murmur_mix(uint32_t h, uint32_t k)
{
  k *= 0xcc9e2d51u;
  k = rotl32(k, 15);
  k *= 0x1b873593u;
  h ^= k;
  h = rotl32(h, 13);
  return h * 5u + 0xe6546b64u;
}
__device__ __forceinline__ uint32_t murmur_fin(uint32_t h, uint32_t len)
{
  h ^= len;
  h ^= h >> 16;
  h *= 0x85ebca6bu;
  h ^= h >> 13;
  h *= 0xc2b2ae35u;
  h ^= h >> 16;
  return h;
}
__device__ __forceinline__ uint32_t murmur3(int32_t v)
{
  return murmur_fin(murmur_mix(0u, static_cast<uint32_t>(v)), 4u);
}
__device__ __forceinline__ uint32_t murmur3(int64_t v)
{
  uint64_t u = static_cast<uint64_t>(v);
  uint32_t h = murmur_mix(0u, static_cast<uint32_t>(u));
  h          = murmur_mix(h, static_cast<uint32_t>(u >> 32));
  return murmur_fin(h, 8u);
}

// compute_gpu_id_from_ext_edge_endpoints_t (graph_utils.cuh:69-82)
template <typename VT>
__global__ void edge_owner_kernel(const VT* majors, const VT* minors, size_t n, uint32_t P,
                                  uint32_t R, uint32_t* dest, uint32_t* iota)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    uint32_t mj = murmur3(majors[i]) % P;
    uint32_t mn = murmur3(minors[i]) % P;
    dest[i]     = (mn / R) * R + (mj % R);
    iota[i]     = static_cast<uint32_t>(i);
  }
}

// compute_gpu_id_from_ext_vertex_t (graph_utils.cuh:45-53); `div` = 1 gives the owner rank,
// `div` = R the owner's col_comm rank = local edge partition index of a major
template <typename VT>
__global__ void vertex_owner_kernel(const VT* v, size_t n, uint32_t P, uint32_t div,
                                    uint32_t* dest, uint32_t* iota)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    dest[i] = (murmur3(v[i]) % P) / div;
    iota[i] = static_cast<uint32_t>(i);
  }
}

template <typename T>
__global__ void gather_kernel(const T* in, const uint32_t* order, size_t n, T* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = in[order[i]];
}

// bounds[p] = first position of the ascending array `sorted` holding a value >= p  (p = 0..P)
__global__ void group_bounds_kernel(const uint32_t* sorted, size_t n, int P, unsigned long long* bounds)
{
  int p = threadIdx.x;
  if (p > P) return;
  size_t lo = 0, hi = n;
  while (lo < hi) {
    size_t mid = (lo + hi) >> 1;
    if (sorted[mid] < static_cast<uint32_t>(p)) lo = mid + 1;
    else hi = mid;
  }
  bounds[p] = lo;
}

template <typename VT>
__device__ __forceinline__ int64_t lower_bound_dev(const VT* a, int64_t n, VT key)
{
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// global major degree of the local vertices: every received major is one edge
template <typename VT>
__global__ void count_majors_kernel(const VT* majors, size_t n, const VT* sorted_local, int64_t nloc,
                                    uint32_t* deg, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    VT m      = majors[i];
    int64_t p = lower_bound_dev(sorted_local, nloc, m);
    if (p >= nloc || sorted_local[p] != m) { *bad = 1; continue; }
    atomicAdd(&deg[p], 1u);
  }
}

// position of every edge endpoint in the gathered label arrays -> sort key
//   key = (position of the major in the concatenated major labels) << 32 | minor offset
template <typename VT>
__global__ void mg_edge_keys_kernel(const VT* majors, const VT* minors, size_t n,
                                    const uint64_t* major_ext, const uint32_t* major_pos,
                                    int64_t n_major, const uint64_t* minor_ext,
                                    const uint32_t* minor_pos, int64_t n_minor, uint64_t* keys,
                                    uint32_t* vals, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    uint64_t mj = static_cast<uint64_t>(majors[i]);
    uint64_t mn = static_cast<uint64_t>(minors[i]);
    int64_t pj  = lower_bound_dev(major_ext, n_major, mj);
    int64_t pn  = lower_bound_dev(minor_ext, n_minor, mn);
    if (pj >= n_major || major_ext[pj] != mj || pn >= n_minor || minor_ext[pn] != mn) {
      *bad    = 1;
      keys[i] = 0;
      vals[i] = static_cast<uint32_t>(i);
      continue;
    }
    keys[i] = (static_cast<uint64_t>(major_pos[pj]) << 32) | minor_pos[pn];
    vals[i] = static_cast<uint32_t>(i);
  }
}

template <typename VT>
__global__ void label_keys_kernel(const VT* labels, size_t n, uint64_t* keys, uint32_t* vals)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    keys[i] = static_cast<uint64_t>(labels[i]);
    vals[i] = static_cast<uint32_t>(i);
  }
}

// row_start[r] = first sorted edge whose row (key >> 32) is >= r, for r = 0..n_rows
__global__ void row_starts_kernel(const uint64_t* keys, size_t ne, int64_t n_rows,
                                  unsigned long long* row_start)
{
  for (size_t e = blockIdx.x * size_t(blockDim.x) + threadIdx.x; e <= ne;
       e += size_t(gridDim.x) * blockDim.x) {
    int64_t prev = (e == 0) ? -1 : static_cast<int64_t>(keys[e - 1] >> 32);
    int64_t cur  = (e == ne) ? n_rows : static_cast<int64_t>(keys[e] >> 32);
    for (int64_t r = prev + 1; r <= cur; ++r) row_start[r] = e;
  }
}

template <typename ET>
__global__ void partition_offsets_kernel(const unsigned long long* row_start, int64_t row_first,
                                         int64_t n_rows, ET* offsets)
{
  unsigned long long base = row_start[row_first];
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r <= n_rows;
       r += int64_t(gridDim.x) * blockDim.x)
    offsets[r] = static_cast<ET>(row_start[row_first + r] - base);
}

// offsets of the fused rectangle: partition i's rows start at row i * slice (slice = the largest
// major range, so that the C partial-sum slices have one size and ONE reduce-scatter combines
// them); the rows behind a partition's last major are empty
template <typename ET>
__global__ void fused_offsets_kernel(const unsigned long long* row_start, const int64_t* piece_first,
                                     int C, int64_t slice, ET* offsets)
{
  const int64_t total = static_cast<int64_t>(C) * slice;
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r <= total;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t i = r == total ? C - 1 : r / slice;
    int64_t k       = r - i * slice;
    const int64_t n = piece_first[i + 1] - piece_first[i];
    if (k > n) k = n;
    offsets[r] = static_cast<ET>(row_start[piece_first[i] + k]);
  }
}

// DCSC (compress_hypersparse_offsets, structure_utils.cuh:109-159): nz[r - hs] = 1 when row r of
// the hypersparse segment has edges; after the scan of nz, the non-empty rows are listed and the
// compressed offsets written
template <typename ET>
__global__ void dcs_flags_kernel(const ET* offsets, int64_t hs, int64_t n_rows, uint32_t* nz)
{
  for (int64_t r = hs + blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < n_rows;
       r += int64_t(gridDim.x) * blockDim.x)
    nz[r - hs] = offsets[r + 1] > offsets[r] ? 1u : 0u;
}
template <typename ET>
__global__ void dcs_fill_kernel(const ET* offsets, const uint32_t* pos, int64_t hs, int64_t n_rows,
                                int64_t major_first, ET* dcs_offsets, int32_t* dcs_nzd)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r <= n_rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    if (r <= hs) dcs_offsets[r] = offsets[r];   // rows before the segment + the segment's start
    if (r < n_rows && r >= hs && offsets[r + 1] > offsets[r]) {
      const uint32_t k       = pos[r - hs];
      dcs_nzd[k]             = static_cast<int32_t>(major_first + r);
      dcs_offsets[hs + k + 1] = offsets[r + 1];
    }
  }
}

__global__ void add_u64_to_double_kernel(const unsigned long long* in, size_t n, double* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<double>(in[i]);
}

// ---- host helpers -------------------------------------------------------------------------
struct shuffle_plan {
  dbuf<uint32_t> order;            // element indices grouped by destination (stable)
  std::vector<size_t> send_counts;
};

// dest[] / iota[] are consumed (sorted in place or in the spare buffers)
shuffle_plan make_plan(b200_handle* h, dbuf<uint32_t>& dest, dbuf<uint32_t>& iota, size_t n, int P)
{
  cudaStream_t s = h->stream;
  shuffle_plan plan;
  plan.send_counts.assign(P, 0);
  if (P == 1) {
    plan.send_counts[0] = n;
    plan.order          = std::move(iota);
    return plan;
  }
  dbuf<uint32_t> k1(n, s), v1(n, s);
  int where = radix_sort<uint32_t>(dest.data(), k1.data(), iota.data(), v1.data(), n,
                                   {{0, bits_needed(static_cast<uint64_t>(P - 1))}}, s);
  const uint32_t* sorted = where ? k1.data() : dest.data();
  dbuf<unsigned long long> d_bounds(P + 1, s);
  group_bounds_kernel<<<1, 64, 0, s>>>(sorted, n, P, d_bounds.data());
  CUDA_TRY(cudaGetLastError());
  std::vector<unsigned long long> b(P + 1);
  CUDA_TRY(cudaMemcpyAsync(b.data(), d_bounds.data(), (P + 1) * 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  for (int p = 0; p < P; ++p) plan.send_counts[p] = static_cast<size_t>(b[p + 1] - b[p]);
  plan.order = where ? std::move(v1) : std::move(iota);
  return plan;
}

template <typename T>
void gather_to(b200_handle* h, const T* in, const uint32_t* order, size_t n, T* out)
{
  if (n == 0) return;
  gather_kernel<T><<<grid_for(n, 4), kThreads, 0, h->stream>>>(in, order, n, out);
  CUDA_TRY(cudaGetLastError());
}

// permute `in` by plan.order and exchange; returns the received elements
template <typename T>
dbuf<T> shuffle(b200_handle* h, ncclComm_t comm, int comm_size, int comm_rank, const T* in, size_t n,
                shuffle_plan const& plan, std::vector<size_t> const& recv_counts)
{
  cudaStream_t s = h->stream;
  dbuf<T> grouped(n, s);
  gather_to<T>(h, in, plan.order.data(), n, grouped.data());
  size_t n_recv = std::accumulate(recv_counts.begin(), recv_counts.end(), size_t{0});
  dbuf<T> out(n_recv, s);
  alltoallv(h, comm, comm_size, comm_rank, grouped.data(), plan.send_counts, out.data(), recv_counts,
            sizeof(T));
  CUDA_TRY(cudaStreamSynchronize(s));
  return out;
}

// ascending unique of n ids given as VT: widen -> radix sort -> compact
template <typename VT>
dbuf<VT> sort_unique(b200_handle* h, const VT* a, const VT* b, size_t n_each, int64_t* n_out)
{
  cudaStream_t s = h->stream;
  size_t n_all   = b ? 2 * n_each : n_each;
  dbuf<VT> out(n_all, s);
  *n_out = 0;
  if (n_all == 0) return out;
  dbuf<uint64_t> k0(n_all, s), k1(n_all, s);
  launch_widen_ids(a, b, n_each, sizeof(VT), k0.data(), s);
  int where = radix_sort<uint64_t>(k0.data(), k1.data(), nullptr, nullptr, n_all,
                                   {{0, static_cast<int>(sizeof(VT) * 8)}}, s);
  *n_out = launch_unique_ids(where ? k1.data() : k0.data(), n_all, sizeof(VT), out.data(), s);
  return out;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// the build
// ------------------------------------------------------------------------------------------
template <typename VT>
static void create_mg_t(b200_handle* h, const VT* src, const VT* dst, const void* weights,
                        size_t ne_in, int weight_type, const VT* vertices, size_t n_given,
                        bool store_transposed, b200_graph* g)
{
  cudaStream_t s = h->stream;
  const int P = h->size, R = h->row_size, C = h->col_size;
  const int row_rank = h->row_rank, col_rank = h->col_rank;
  const VT* in_majors = store_transposed ? dst : src;
  const VT* in_minors = store_transposed ? src : dst;
  const size_t wsize  = weights ? static_cast<size_t>(dtype_size(weight_type)) : 0;
  B200_EXPECTS(ne_in < (size_t{1} << 32), kInvalidInput, "too many edges on one rank");

  dbuf<int> d_flag(1, s);
  dbuf<unsigned long long> d_max(1, s);
  int h_flag = 0;
  CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
  CUDA_TRY(cudaMemsetAsync(d_max.data(), 0, 8, s));
  if (ne_in) {
    max_id_kernel<VT><<<std::min(grid_for(ne_in, 8), 4096u), kThreads, 0, s>>>(
      src, dst, ne_in, d_max.data(), d_flag.data());
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(!h_flag, kInvalidInput, "Invalid input argument: negative vertex id.");

  // ---- 1. edges to their 2D owner (shuffle_wrappers.cu:41-174) ----
  dbuf<VT> e_major, e_minor;
  dbuf<char> e_w;
  size_t ne = 0;
  {
    dbuf<uint32_t> dest(ne_in, s), iota(ne_in, s);
    if (ne_in) {
      edge_owner_kernel<VT><<<grid_for(ne_in, 4), kThreads, 0, s>>>(
        in_majors, in_minors, ne_in, static_cast<uint32_t>(P), static_cast<uint32_t>(R), dest.data(),
        iota.data());
      CUDA_TRY(cudaGetLastError());
    }
    shuffle_plan plan = make_plan(h, dest, iota, ne_in, P);
    std::vector<size_t> recv = exchange_counts(h, h->world, P, h->rank, plan.send_counts);
    e_major = shuffle<VT>(h, h->world, P, h->rank, in_majors, ne_in, plan, recv);
    e_minor = shuffle<VT>(h, h->world, P, h->rank, in_minors, ne_in, plan, recv);
    ne      = e_major.size();
    if (weights) {
      if (wsize == 4) {
        dbuf<float> w = shuffle<float>(h, h->world, P, h->rank, static_cast<const float*>(weights),
                                       ne_in, plan, recv);
        e_w.p = reinterpret_cast<char*>(w.p); e_w.n = w.n * 4; e_w.st = s; w.p = nullptr; w.n = 0;
      } else {
        dbuf<double> w = shuffle<double>(h, h->world, P, h->rank,
                                         static_cast<const double*>(weights), ne_in, plan, recv);
        e_w.p = reinterpret_cast<char*>(w.p); e_w.n = w.n * 8; e_w.st = s; w.p = nullptr; w.n = 0;
      }
    }
  }
  B200_EXPECTS(ne < (size_t{1} << 32), kInvalidInput, "too many edges on one rank");

  // ---- 2. local vertex list: unique endpoints (or the given list) sent to their hash owner
  //         (renumber_edgelist_impl.cuh:108-217) ----
  dbuf<VT> sorted_local;
  int64_t nloc = 0;
  {
    int64_t n_u = 0;
    dbuf<VT> uniq = (vertices && n_given) ? sort_unique<VT>(h, vertices, nullptr, n_given, &n_u)
                                          : sort_unique<VT>(h, e_major.data(), e_minor.data(), ne, &n_u);
    dbuf<uint32_t> dest(static_cast<size_t>(n_u), s), iota(static_cast<size_t>(n_u), s);
    if (n_u) {
      vertex_owner_kernel<VT><<<grid_for(n_u, 4), kThreads, 0, s>>>(
        uniq.data(), static_cast<size_t>(n_u), static_cast<uint32_t>(P), 1u, dest.data(), iota.data());
      CUDA_TRY(cudaGetLastError());
    }
    shuffle_plan plan = make_plan(h, dest, iota, static_cast<size_t>(n_u), P);
    std::vector<size_t> recv = exchange_counts(h, h->world, P, h->rank, plan.send_counts);
    dbuf<VT> mine = shuffle<VT>(h, h->world, P, h->rank, uniq.data(), static_cast<size_t>(n_u), plan, recv);
    sorted_local  = sort_unique<VT>(h, mine.data(), nullptr, mine.size(), &nloc);
  }
  B200_EXPECTS(nloc < (int64_t{1} << 31), kInvalidInput, "too many vertices on one rank");

  // ---- 3. group the local edges by edge partition (= col_comm rank of the major's owner) and
  //         count global major degrees on the owners (renumber_edgelist_impl.cuh:219-382) ----
  std::vector<size_t> part_edges(C, 0);
  dbuf<uint32_t> deg(static_cast<size_t>(nloc), s);
  CUDA_TRY(cudaMemsetAsync(deg.data(), 0, std::max<size_t>(static_cast<size_t>(nloc), 1) * 4, s));
  {
    dbuf<uint32_t> dest(ne, s), iota(ne, s);
    if (ne) {
      vertex_owner_kernel<VT><<<grid_for(ne, 4), kThreads, 0, s>>>(
        e_major.data(), ne, static_cast<uint32_t>(P), static_cast<uint32_t>(R), dest.data(),
        iota.data());
      CUDA_TRY(cudaGetLastError());
    }
    shuffle_plan plan = make_plan(h, dest, iota, ne, C);
    part_edges        = plan.send_counts;
    // keep the edges grouped by partition from here on
    dbuf<VT> gm(ne, s), gn(ne, s);
    gather_to<VT>(h, e_major.data(), plan.order.data(), ne, gm.data());
    gather_to<VT>(h, e_minor.data(), plan.order.data(), ne, gn.data());
    if (weights) {
      dbuf<char> gw(ne * wsize, s);
      if (wsize == 4)
        gather_to<float>(h, reinterpret_cast<const float*>(e_w.data()), plan.order.data(), ne,
                         reinterpret_cast<float*>(gw.data()));
      else
        gather_to<double>(h, reinterpret_cast<const double*>(e_w.data()), plan.order.data(), ne,
                          reinterpret_cast<double*>(gw.data()));
      CUDA_TRY(cudaStreamSynchronize(s));
      e_w = std::move(gw);
    }
    CUDA_TRY(cudaStreamSynchronize(s));
    e_major = std::move(gm);
    e_minor = std::move(gn);
    // majors of partition i go to col_comm rank i, which owns them
    std::vector<size_t> recv = exchange_counts(h, h->col, C, col_rank, part_edges);
    size_t n_recv = std::accumulate(recv.begin(), recv.end(), size_t{0});
    dbuf<VT> owned(n_recv, s);
    alltoallv(h, h->col, C, col_rank, e_major.data(), part_edges, owned.data(), recv, sizeof(VT));
    CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
    if (n_recv) {
      count_majors_kernel<VT><<<grid_for(n_recv, 4), kThreads, 0, s>>>(
        owned.data(), n_recv, sorted_local.data(), nloc, deg.data(), d_flag.data());
      CUDA_TRY(cudaGetLastError());
    }
    CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  {
    // every rank must leave together, whoever saw the bad endpoint
    std::vector<int64_t> flags = allgather_host(h, h->world, P, {static_cast<int64_t>(h_flag)});
    for (int64_t f : flags)
      B200_EXPECTS(!f, kInvalidInput,
                   "Invalid input argument: edge endpoints are not in the vertex list.");
  }

  // ---- 4. stable sort by descending global degree + segment offsets with the thresholds
  //         scaled by col_comm_size (renumber_edgelist_impl.cuh:384-449) ----
  g->renumber_map.alloc(static_cast<size_t>(nloc) * sizeof(VT), s);
  VT* labels = reinterpret_cast<VT*>(g->renumber_map.data());
  std::vector<int64_t> thr;
  {
    int64_t mid = kMidDegreeThreshold * C, low = kLowDegreeThreshold * C;
    int64_t hyp = static_cast<int64_t>(C * kHypersparseRatio);
    if (hyp > 0) thr = {mid, low, hyp, std::min<int64_t>(hyp, 1)};
    else thr = {mid, low, 1};
  }
  std::vector<int64_t> my_seg(thr.size() + 2, 0);
  my_seg.back() = nloc;
  if (nloc > 0) {
    dbuf<uint32_t> d_maxdeg(1, s);
    CUDA_TRY(cudaMemsetAsync(d_maxdeg.data(), 0, 4, s));
    max_u32_kernel<<<std::min(grid_for(nloc, 8), 4096u), kThreads, 0, s>>>(deg.data(), nloc,
                                                                        d_maxdeg.data());
    CUDA_TRY(cudaGetLastError());
    uint32_t maxdeg = 0;
    CUDA_TRY(cudaMemcpyAsync(&maxdeg, d_maxdeg.data(), 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    dbuf<uint32_t> k0(nloc, s), k1(nloc, s), v0(nloc, s), v1(nloc, s);
    degree_keys_kernel<<<grid_for(nloc, 4), kThreads, 0, s>>>(deg.data(), nloc, maxdeg, k0.data(),
                                                            v0.data());
    CUDA_TRY(cudaGetLastError());
    int where = radix_sort<uint32_t>(k0.data(), k1.data(), v0.data(), v1.data(), nloc,
                                     {{0, bits_needed(maxdeg)}}, s);
    dbuf<uint32_t> new_of_sorted(nloc, s), sorted_deg(nloc, s);
    finish_renumber_kernel<VT><<<grid_for(nloc, 4), kThreads, 0, s>>>(
      where ? k1.data() : k0.data(), where ? v1.data() : v0.data(), nloc, maxdeg,
      sorted_local.data(), 0, labels, new_of_sorted.data(), sorted_deg.data());
    CUDA_TRY(cudaGetLastError());
    dbuf<int64_t> d_thr(thr.size(), s), d_seg(thr.size(), s);
    CUDA_TRY(cudaMemcpyAsync(d_thr.data(), thr.data(), thr.size() * 8, cudaMemcpyHostToDevice, s));
    segment_bounds_kernel<<<1, 32, 0, s>>>(sorted_deg.data(), nloc, d_thr.data(),
                                           static_cast<int>(thr.size()), d_seg.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(my_seg.data() + 1, d_seg.data(), thr.size() * 8,
                             cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }

  // ---- 5. vertex partition ranges and every rank's segment offsets (:758-771, :907-914) ----
  const int nseg = static_cast<int>(my_seg.size());
  std::vector<int64_t> mine_meta(my_seg);
  mine_meta.push_back(static_cast<int64_t>(ne));
  std::vector<int64_t> all_meta = allgather_host(h, h->world, P, mine_meta);
  std::vector<int64_t> vpo(P + 1, 0);
  int64_t total_edges = 0;
  for (int p = 0; p < P; ++p) {
    vpo[p + 1] = vpo[p] + all_meta[p * (nseg + 1) + nseg - 1];
    total_edges += all_meta[p * (nseg + 1) + nseg];
  }
  B200_EXPECTS(vpo[P] < (int64_t{1} << 31), kInvalidInput,
               "number of vertices must fit 32-bit internal ids");

  // ---- 6. labels of the vertex partitions this rank's edges touch: majors come from the
  //         col_comm peers, minors from the row_comm peers (:793-905); one sorted
  //         (external id -> position) table each replaces the reference's cuco maps ----
  std::vector<size_t> major_counts(C), minor_counts(R);
  std::vector<int64_t> major_piece_first(C + 1, 0);
  for (int i = 0; i < C; ++i) {
    major_counts[i]          = static_cast<size_t>(vpo[R * i + row_rank + 1] - vpo[R * i + row_rank]);
    major_piece_first[i + 1] = major_piece_first[i] + static_cast<int64_t>(major_counts[i]);
  }
  for (int j = 0; j < R; ++j)
    minor_counts[j] = static_cast<size_t>(vpo[col_rank * R + j + 1] - vpo[col_rank * R + j]);
  const int64_t n_major     = major_piece_first[C];
  const int64_t minor_first = vpo[col_rank * R];
  const int64_t n_minor     = vpo[(col_rank + 1) * R] - minor_first;

  auto build_table = [&](ncclComm_t comm, int csize, int crank, std::vector<size_t> const& counts,
                         int64_t n_total, dbuf<uint64_t>& ext, dbuf<uint32_t>& pos) {
    dbuf<VT> gathered(static_cast<size_t>(n_total), s);
    allgatherv(s, comm, csize, crank, labels, gathered.data(), counts, sizeof(VT));
    dbuf<uint64_t> k0(static_cast<size_t>(n_total), s), k1(static_cast<size_t>(n_total), s);
    dbuf<uint32_t> v0(static_cast<size_t>(n_total), s), v1(static_cast<size_t>(n_total), s);
    if (n_total) {
      label_keys_kernel<VT><<<grid_for(n_total, 4), kThreads, 0, s>>>(
        gathered.data(), static_cast<size_t>(n_total), k0.data(), v0.data());
      CUDA_TRY(cudaGetLastError());
    }
    int where = radix_sort<uint64_t>(k0.data(), k1.data(), v0.data(), v1.data(),
                                     static_cast<size_t>(n_total),
                                     {{0, static_cast<int>(sizeof(VT) * 8)}}, s);
    CUDA_TRY(cudaStreamSynchronize(s));
    ext = where ? std::move(k1) : std::move(k0);
    pos = where ? std::move(v1) : std::move(v0);
  };
  dbuf<uint64_t> major_ext, minor_ext;
  dbuf<uint32_t> major_pos, minor_pos;
  build_table(h->col, C, col_rank, major_counts, n_major, major_ext, major_pos);
  build_table(h->row, R, row_rank, minor_counts, n_minor, minor_ext, minor_pos);

  // ---- 7. one stable sort lays out every local partition: rows in (partition, local row)
  //         order, neighbours ascending, duplicates in arrival order ----
  dbuf<uint64_t> k0(ne, s), k1(ne, s);
  dbuf<uint32_t> v0(ne, s), v1(ne, s);
  CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
  if (ne) {
    mg_edge_keys_kernel<VT><<<grid_for(ne, 4), kThreads, 0, s>>>(
      e_major.data(), e_minor.data(), ne, major_ext.data(), major_pos.data(), n_major,
      minor_ext.data(), minor_pos.data(), n_minor, k0.data(), v0.data(), d_flag.data());
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(!h_flag, kUnknownError, "multi-GPU renumbering: an edge endpoint has no label");
  e_major.release();
  e_minor.release();
  major_ext.release(); major_pos.release(); minor_ext.release(); minor_pos.release();
  int where = radix_sort<uint64_t>(
    k0.data(), k1.data(), v0.data(), v1.data(), ne,
    {{0, bits_needed(static_cast<uint64_t>(std::max<int64_t>(n_minor - 1, 0)))},
     {32, 32 + bits_needed(static_cast<uint64_t>(std::max<int64_t>(n_major - 1, 0)))}},
    s);
  const uint64_t* ks = where ? k1.data() : k0.data();
  const uint32_t* vs = where ? v1.data() : v0.data();

  dbuf<unsigned long long> row_start(static_cast<size_t>(n_major) + 1, s);
  row_starts_kernel<<<grid_for(ne + 1, 4), kThreads, 0, s>>>(ks, ne, n_major, row_start.data());
  CUDA_TRY(cudaGetLastError());
  std::vector<unsigned long long> piece_edge_first(C + 1, 0);
  for (int i = 0; i <= C; ++i)
    CUDA_TRY(cudaMemcpyAsync(&piece_edge_first[i], row_start.data() + major_piece_first[i], 8,
                             cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));

  // ---- 8. the edge partitions ----
  g->parts.clear();
  g->parts.resize(C);
  g->partition_segment_offsets.assign(C, {});
  for (int i = 0; i < C; ++i) {
    edge_partition& ep = g->parts[i];
    const size_t e0 = piece_edge_first[i], e1 = piece_edge_first[i + 1];
    const size_t n  = e1 - e0;
    B200_EXPECTS(n == part_edges[i], kUnknownError, "multi-GPU build: partition sizes disagree");
    ep.num_rows    = static_cast<int64_t>(major_counts[i]);
    ep.num_edges   = static_cast<int64_t>(n);
    ep.major_first = vpo[R * i + row_rank];
    ep.minor_first = minor_first;
    ep.num_minors  = n_minor;
    ep.minor_piece_offsets.assign(1, 0);
    for (int j = 0; j < R; ++j)
      ep.minor_piece_offsets.push_back(ep.minor_piece_offsets.back() +
                                       static_cast<int64_t>(minor_counts[j]));
    ep.offsets64   = n >= static_cast<size_t>(0x7fffffff);
    ep.indices.alloc(n, s, 16);
    if (weights) ep.weights.alloc(n * wsize, s, 32);
    if (n) {
      if (!weights)
        finish_edges_kernel<float><<<grid_for(n, 4), kThreads, 0, s>>>(ks + e0, vs + e0, n, nullptr,
                                                                     ep.indices.data(), nullptr);
      else if (wsize == 4)
        finish_edges_kernel<float><<<grid_for(n, 4), kThreads, 0, s>>>(
          ks + e0, vs + e0, n, reinterpret_cast<const float*>(e_w.data()), ep.indices.data(),
          reinterpret_cast<float*>(ep.weights.data()));
      else
        finish_edges_kernel<double><<<grid_for(n, 4), kThreads, 0, s>>>(
          ks + e0, vs + e0, n, reinterpret_cast<const double*>(e_w.data()), ep.indices.data(),
          reinterpret_cast<double*>(ep.weights.data()));
      CUDA_TRY(cudaGetLastError());
    }
    const size_t nr = static_cast<size_t>(ep.num_rows);
    if (ep.offsets64) {
      ep.offsets.alloc((nr + 1) * 8, s, 64);
      partition_offsets_kernel<int64_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
        row_start.data(), major_piece_first[i], ep.num_rows,
        reinterpret_cast<int64_t*>(ep.offsets.data()));
    } else {
      ep.offsets.alloc((nr + 1) * 4, s, 64);
      partition_offsets_kernel<int32_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
        row_start.data(), major_piece_first[i], ep.num_rows,
        reinterpret_cast<int32_t*>(ep.offsets.data()));
    }
    CUDA_TRY(cudaGetLastError());
    // segment offsets of the vertex partition the majors belong to (aggregate_segment_offsets,
    // renumber_edgelist_impl.cuh:907-908); local degrees are not monotone, so the kernel
    // schedule is packed from the offsets themselves
    const int owner = R * i + row_rank;
    std::vector<int64_t> seg(all_meta.begin() + owner * (nseg + 1),
                             all_meta.begin() + owner * (nseg + 1) + nseg);
    g->partition_segment_offsets[i] = seg;
    // 8b. DCSC form of the hypersparse segment, as the reference holds it (views and parity
    //     tests; the kernels' layout is built from the uncompressed offsets)
    ep.hypersparse_first = -1;
    if (nseg == 6) {
      const int64_t hs = seg[3];
      const int64_t nh = ep.num_rows - hs;
      ep.hypersparse_first = hs;
      dbuf<uint32_t> nz(static_cast<size_t>(nh) + 1, s);
      uint32_t cnt = 0;
      if (nh > 0) {
        if (ep.offsets64)
          dcs_flags_kernel<int64_t><<<grid_for(static_cast<size_t>(nh), 4), kThreads, 0, s>>>(
            reinterpret_cast<const int64_t*>(ep.offsets.data()), hs, ep.num_rows, nz.data());
        else
          dcs_flags_kernel<int32_t><<<grid_for(static_cast<size_t>(nh), 4), kThreads, 0, s>>>(
            reinterpret_cast<const int32_t*>(ep.offsets.data()), hs, ep.num_rows, nz.data());
        CUDA_TRY(cudaGetLastError());
        exclusive_scan<uint32_t, uint32_t>(nz.data(), nz.data(), static_cast<size_t>(nh), s, true);
        CUDA_TRY(cudaMemcpyAsync(&cnt, nz.data() + nh, 4, cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
      }
      ep.dcs_nzd_count = cnt;
      const size_t esz = ep.offsets64 ? 8 : 4;
      ep.dcs_offsets.alloc((static_cast<size_t>(hs) + cnt + 1) * esz, s, 64);
      ep.dcs_nzd_vertices.alloc(std::max<size_t>(cnt, 1), s);
      if (ep.offsets64)
        dcs_fill_kernel<int64_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
          reinterpret_cast<const int64_t*>(ep.offsets.data()), nz.data(), hs, ep.num_rows, ep.major_first,
          reinterpret_cast<int64_t*>(ep.dcs_offsets.data()), ep.dcs_nzd_vertices.data());
      else
        dcs_fill_kernel<int32_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
          reinterpret_cast<const int32_t*>(ep.offsets.data()), nz.data(), hs, ep.num_rows, ep.major_first,
          reinterpret_cast<int32_t*>(ep.dcs_offsets.data()), ep.dcs_nzd_vertices.data());
      CUDA_TRY(cudaGetLastError());
      CUDA_TRY(cudaStreamSynchronize(s));   // nz is freed at the end of this scope
    }
    ep.has_segments     = true;
    ep.generic_schedule = true;
    ep.seg[0] = 0; ep.seg[1] = seg[1]; ep.seg[2] = seg[2]; ep.seg[3] = seg[nseg - 2];
    ep.seg[4] = ep.num_rows;
  }
  // ---- 9. (default since round 2; B200_MG_FUSED=0 turns it off) the same edges once more as a
  //         single rectangle over the concatenated major ranges: one launch train and one
  //         grouped reduction per iteration instead of C of each; see mg.cu ----
  g->fused.clear();
  g->fused_row_first.assign(major_piece_first.begin(), major_piece_first.end());
  g->fused_slice = 0;
  const char* want_fused = std::getenv("B200_MG_FUSED");
  if (C > 1 && (want_fused ? std::atoi(want_fused) != 0 : true)) {
    int64_t slice = 0;
    for (int i = 0; i < C; ++i) slice = std::max<int64_t>(slice, static_cast<int64_t>(major_counts[i]));
    slice = (slice + 31) & ~int64_t{31};
    g->fused_slice = slice;
    for (int i = 0; i <= C; ++i) g->fused_row_first[i] = i * slice;
    const int64_t n_rows_f = slice * C;
    g->fused.resize(1);
    edge_partition& fp = g->fused[0];
    fp.num_rows    = n_rows_f;
    fp.num_edges   = static_cast<int64_t>(ne);
    fp.major_first = -1;  // not one global range
    fp.minor_first = minor_first;
    fp.num_minors  = n_minor;
    fp.minor_piece_offsets = g->parts[0].minor_piece_offsets;
    fp.offsets64   = ne >= static_cast<size_t>(0x7fffffff);
    fp.indices.alloc(ne, s, 16);
    if (weights) fp.weights.alloc(ne * wsize, s, 32);
    if (ne) {
      if (!weights)
        finish_edges_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(ks, vs, ne, nullptr,
                                                                      fp.indices.data(), nullptr);
      else if (wsize == 4)
        finish_edges_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ks, vs, ne, reinterpret_cast<const float*>(e_w.data()), fp.indices.data(),
          reinterpret_cast<float*>(fp.weights.data()));
      else
        finish_edges_kernel<double><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ks, vs, ne, reinterpret_cast<const double*>(e_w.data()), fp.indices.data(),
          reinterpret_cast<double*>(fp.weights.data()));
      CUDA_TRY(cudaGetLastError());
    }
    const size_t nr = static_cast<size_t>(n_rows_f);
    dbuf<int64_t> d_piece(static_cast<size_t>(C) + 1, s);
    CUDA_TRY(cudaMemcpyAsync(d_piece.data(), major_piece_first.data(), (static_cast<size_t>(C) + 1) * 8,
                             cudaMemcpyHostToDevice, s));
    if (fp.offsets64) {
      fp.offsets.alloc((nr + 1) * 8, s, 64);
      fused_offsets_kernel<int64_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
        row_start.data(), d_piece.data(), C, slice, reinterpret_cast<int64_t*>(fp.offsets.data()));
    } else {
      fp.offsets.alloc((nr + 1) * 4, s, 64);
      fused_offsets_kernel<int32_t><<<grid_for(nr + 1, 4), kThreads, 0, s>>>(
        row_start.data(), d_piece.data(), C, slice, reinterpret_cast<int32_t*>(fp.offsets.data()));
    }
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));   // d_piece is freed at the end of this scope
    fp.has_segments     = true;
    fp.generic_schedule = true;
    fp.seg[4]           = n_rows_f;
  }
  CUDA_TRY(cudaStreamSynchronize(s));

  g->num_vertices             = vpo[P];
  g->num_edges                = total_edges;
  g->local_first              = vpo[h->rank];
  g->local_last               = vpo[h->rank + 1];
  g->vertex_partition_offsets = vpo;
  g->renumbered               = true;
  g->num_segments             = nseg - 1;
  for (int i = 0; i < 8; ++i) g->segment_offsets[i] = i < nseg ? my_seg[i] : 0;
  bool any64 = false;
  for (auto const& ep : g->parts) any64 = any64 || ep.offsets64;
  g->edge_type = any64 ? B200_INT64 : B200_INT32;
}

void graph_create_mg(b200_handle* h, const void* src, const void* dst, const void* weights,
                     size_t ne, int vertex_type, int weight_type, const void* local_vertices,
                     size_t n_local_vertices, bool store_transposed, b200_graph* g)
{
  B200_EXPECTS(h->world != nullptr, kInvalidHandle, "multi-GPU graph needs an MG handle");
  // Weighted or not is a property of the whole graph (the reference passes an optional vector,
  // present on every rank even when empty, graph_functions.hpp:463-478).  A rank that
  // contributes no edges may hold a null pointer, so the ranks agree first: every rank must make
  // the same sequence of exchanges.
  dbuf<char> no_weights;
  {
    int64_t mine = ne == 0 ? 0 : (weights ? 2 + weight_type : 1);
    std::vector<int64_t> all = allgather_host(h, h->world, h->size, {mine});
    int64_t agreed = 0;
    bool mixed     = false;
    for (int64_t v : all) {
      if (v == 0) continue;
      if (agreed == 0) agreed = v;
      else if (agreed != v) mixed = true;
    }
    B200_EXPECTS(!mixed, kInvalidInput,
                 "Invalid input argument: every rank must pass weights of one type, or none.");
    if (agreed >= 2) {
      weight_type = static_cast<int>(agreed - 2);
      if (!weights) {
        no_weights.alloc(8, h->stream);
        weights = no_weights.data();
      }
    } else if (agreed == 1) {
      weights = nullptr;
    }
  }
  g->vertex_type      = vertex_type;
  g->weight_type      = weights ? weight_type : B200_FLOAT32;
  g->weighted         = weights != nullptr;
  g->store_transposed = store_transposed;
  g->multi_gpu        = true;
  if (vertex_type == B200_INT32)
    create_mg_t<int32_t>(h, static_cast<const int32_t*>(src), static_cast<const int32_t*>(dst),
                         weights, ne, weight_type, static_cast<const int32_t*>(local_vertices),
                         n_local_vertices, store_transposed, g);
  else
    create_mg_t<int64_t>(h, static_cast<const int64_t*>(src), static_cast<const int64_t*>(dst),
                         weights, ne, weight_type, static_cast<const int64_t*>(local_vertices),
                         n_local_vertices, store_transposed, g);
}

// ------------------------------------------------------------------------------------------
// out-weight sums of the local vertices (graph_view_impl.cuh:238-265 on a 2D layout): every
// rank of a row communicator histograms the same minor range; rank j of the row reduces piece j
// ------------------------------------------------------------------------------------------
void graph_out_weight_sums_mg(b200_handle* h, b200_graph* g)
{
  if (g->have_out_weight_sums) return;
  cudaStream_t s = h->stream;
  const int R = h->row_size, col_rank = h->col_rank;
  const size_t n_minor = g->parts.empty() ? 0 : static_cast<size_t>(g->parts[0].num_minors);
  const size_t nloc    = static_cast<size_t>(g->local_last - g->local_first);
  g->out_weight_sums.alloc(nloc, s);
  dbuf<double> sums(n_minor, s);
  if (g->weighted) {
    CUDA_TRY(cudaMemsetAsync(sums.data(), 0, std::max<size_t>(n_minor, 1) * 8, s));
    for (auto& ep : g->parts) {
      size_t ne = static_cast<size_t>(ep.num_edges);
      if (!ne) continue;
      if (g->weight_type == B200_FLOAT32)
        out_weight_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ep.indices.data(), reinterpret_cast<const float*>(ep.weights.data()), ne, sums.data());
      else
        out_weight_kernel<double><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ep.indices.data(), reinterpret_cast<const double*>(ep.weights.data()), ne, sums.data());
      CUDA_TRY(cudaGetLastError());
    }
  } else {
    dbuf<unsigned long long> cnt(n_minor, s);
    CUDA_TRY(cudaMemsetAsync(cnt.data(), 0, std::max<size_t>(n_minor, 1) * 8, s));
    for (auto& ep : g->parts) {
      size_t ne = static_cast<size_t>(ep.num_edges);
      if (!ne) continue;
      out_degree_kernel<<<grid_for(ne, 4), kThreads, 0, s>>>(ep.indices.data(), ne, cnt.data());
      CUDA_TRY(cudaGetLastError());
    }
    if (n_minor) {
      add_u64_to_double_kernel<<<grid_for(n_minor, 4), kThreads, 0, s>>>(cnt.data(), n_minor,
                                                                        sums.data());
      CUDA_TRY(cudaGetLastError());
    }
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  // piece j of the minor range is the vertex partition of row_comm rank j
  auto const& vpo = g->vertex_partition_offsets;
  if (R == 1) {
    if (nloc)
      CUDA_TRY(cudaMemcpyAsync(g->out_weight_sums.data(), sums.data(), nloc * 8,
                               cudaMemcpyDeviceToDevice, s));
  } else {
    NCCL_TRY(ncclGroupStart());
    for (int j = 0; j < R; ++j) {
      size_t first = static_cast<size_t>(vpo[col_rank * R + j] - vpo[col_rank * R]);
      size_t cnt   = static_cast<size_t>(vpo[col_rank * R + j + 1] - vpo[col_rank * R + j]);
      if (cnt == 0) continue;
      NCCL_TRY(ncclReduce(sums.data() + first, g->out_weight_sums.data(), cnt, ncclDouble, ncclSum, j,
                          h->row, s));
    }
    NCCL_TRY(ncclGroupEnd());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  g->have_out_weight_sums = true;
}

}  // namespace b200
