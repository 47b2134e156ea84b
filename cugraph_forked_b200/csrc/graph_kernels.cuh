// Kernels of the graph build shared by graph.cu (single GPU) and graph_mg.cu (multi-GPU).
// Included inside one translation unit each; everything lives in an anonymous namespace.
#pragma once

#include "graph.cuh"

namespace b200 {

namespace {

constexpr int kThreads = 256;
inline unsigned grid_for(size_t n, int per_thread = 1)
{
  size_t b = (n + size_t(kThreads) * per_thread - 1) / (size_t(kThreads) * per_thread);
  if (b == 0) b = 1;
  if (b > 0x7fffffffu) b = 0x7fffffffu;
  return static_cast<unsigned>(b);
}

template <typename VT>
__global__ void max_id_kernel(const VT* a, const VT* b, size_t n, unsigned long long* out,
                              int* negative)
{
  unsigned long long m = 0;
  int neg              = 0;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    VT x = a[i];
    if (x < 0) neg = 1;
    else if (static_cast<unsigned long long>(x) > m) m = static_cast<unsigned long long>(x);
    if (b) {
      VT y = b[i];
      if (y < 0) neg = 1;
      else if (static_cast<unsigned long long>(y) > m) m = static_cast<unsigned long long>(y);
    }
  }
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) atomicMax(out, m);
  if (neg) *negative = 1;
}

template <typename VT>
__global__ void mark_present_kernel(const VT* a, const VT* b, size_t n, uint32_t* flags)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    flags[a[i]] = 1u;
    if (b) flags[b[i]] = 1u;
  }
}

// flags (0/1) + their exclusive scan -> compacted sorted vertex list
template <typename VT>
__global__ void compact_present_kernel(const uint32_t* flags, const uint32_t* pos, size_t n,
                                       VT* sorted_vertices)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    if (flags[i]) sorted_vertices[pos[i]] = static_cast<VT>(i);
  }
}

template <typename VT>
__global__ void is_identity_kernel(const VT* v, size_t n, int* not_identity)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    if (v[i] != static_cast<VT>(i)) *not_identity = 1;
  }
}

// position of an external id in the sorted unique vertex list
template <typename VT>
struct vertex_lookup {
  int mode;  // 0: identity, 1: direct map (pos[id]), 2: binary search
  const uint32_t* pos;
  const VT* sorted;
  int64_t n;
  __device__ __forceinline__ int64_t operator()(VT id) const
  {
    if (mode == 0) return static_cast<int64_t>(id);
    if (mode == 1) return static_cast<int64_t>(pos[id]);
    int64_t lo = 0, hi = n;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (sorted[mid] < id) lo = mid + 1;
      else hi = mid;
    }
    return lo;
  }
};

template <typename VT>
__global__ void count_degree_kernel(const VT* majors, size_t n, vertex_lookup<VT> lk,
                                    uint32_t* deg, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    int64_t p = lk(majors[i]);
    if (p < 0 || p >= lk.n || (lk.mode == 2 && lk.sorted[p] != majors[i])) { *bad = 1; continue; }
    atomicAdd(&deg[p], 1u);
  }
}

__global__ void max_u32_kernel(const uint32_t* a, size_t n, uint32_t* out)
{
  uint32_t m = 0;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    m = max(m, a[i]);
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) atomicMax(out, m);
}

__global__ void degree_keys_kernel(const uint32_t* deg, size_t n, uint32_t maxdeg, uint32_t* keys,
                                   uint32_t* vals)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    keys[i] = maxdeg - deg[i];  // ascending key == descending degree; LSD radix sort is stable
    vals[i] = static_cast<uint32_t>(i);
  }
}

template <typename VT>
__global__ void finish_renumber_kernel(const uint32_t* sorted_keys, const uint32_t* sorted_vals,
                                       size_t n, uint32_t maxdeg, const VT* sorted_vertices,
                                       int identity, VT* labels, uint32_t* new_of_sorted,
                                       uint32_t* sorted_deg)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    uint32_t v      = sorted_vals[i];
    labels[i]       = identity ? static_cast<VT>(v) : sorted_vertices[v];
    new_of_sorted[v] = static_cast<uint32_t>(i);
    sorted_deg[i]   = maxdeg - sorted_keys[i];
  }
}

// number of leading elements of the descending array `deg` that are >= t
// (thrust::upper_bound(..., greater) of renumber_edgelist_impl.cuh:436-442)
__global__ void segment_bounds_kernel(const uint32_t* deg, int64_t n, const int64_t* thresholds,
                                      int nt, int64_t* out)
{
  int k = threadIdx.x;
  if (k >= nt) return;
  int64_t t  = thresholds[k];
  int64_t lo = 0, hi = n;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (static_cast<int64_t>(deg[mid]) >= t) lo = mid + 1;
    else hi = mid;
  }
  out[k] = lo;
}

template <typename VT>
__global__ void edge_keys_kernel(const VT* majors, const VT* minors, size_t n,
                                 vertex_lookup<VT> lk, const uint32_t* new_of_sorted,
                                 uint64_t* keys, uint32_t* vals, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    int64_t pj = lk(majors[i]);
    int64_t pn = lk(minors[i]);
    if (pn < 0 || pn >= lk.n || (lk.mode == 2 && lk.sorted[pn] != minors[i])) {
      *bad = 1;
      pn   = 0;
    }
    uint32_t mj = new_of_sorted ? new_of_sorted[pj] : static_cast<uint32_t>(pj);
    uint32_t mn = new_of_sorted ? new_of_sorted[pn] : static_cast<uint32_t>(pn);
    keys[i]     = (static_cast<uint64_t>(mj) << 32) | mn;
    vals[i]     = static_cast<uint32_t>(i);
  }
}

template <typename WT>
__global__ void finish_edges_kernel(const uint64_t* keys, const uint32_t* vals, size_t n,
                                    const WT* w_in, int32_t* indices, WT* w_out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    indices[i] = static_cast<int32_t>(keys[i] & 0xffffffffu);
    if (w_in) w_out[i] = w_in[vals[i]];
  }
}

template <typename ET>
__global__ void narrow_offsets_kernel(const uint64_t* in, size_t n, ET* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<ET>(in[i]);
}

template <typename VT>
__global__ void iota_kernel(VT* out, size_t n, VT first)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = first + static_cast<VT>(i);
}

template <typename WT>
__global__ void out_weight_kernel(const int32_t* indices, const WT* w, size_t n, double* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    atomicAdd(&out[indices[i]], static_cast<double>(w[i]));
}
__global__ void out_degree_kernel(const int32_t* indices, size_t n, unsigned long long* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    atomicAdd(&out[indices[i]], 1ull);
}
__global__ void u64_to_double_kernel(const unsigned long long* in, size_t n, double* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<double>(in[i]);
}

template <typename ET>
__global__ void in_degree_kernel(const ET* offsets, size_t n, ET* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = offsets[i + 1] - offsets[i];
}

// ---- tile schedule ---------------------------------------------------------------------
// kind-0 tiles: window k of T edges over the edge range of rows [row_lo, row_hi) owns the
// rows whose FIRST edge lies in the window.
template <typename ET>
__global__ void row_tiles_kernel(const ET* offsets, int64_t row_lo, int64_t row_hi,
                                 int64_t edge_lo, int T, int64_t ntiles, tile_t* tiles)
{
  int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (k >= ntiles) return;
  auto lower = [&](int64_t target) {
    int64_t lo = row_lo, hi = row_hi;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (static_cast<int64_t>(offsets[mid]) < target) lo = mid + 1;
      else hi = mid;
    }
    return lo;
  };
  int64_t r0 = lower(edge_lo + k * T);
  int64_t r1 = (k + 1 == ntiles) ? row_hi : lower(edge_lo + (k + 1) * T);
  tile_t t;
  t.edge_first = static_cast<int64_t>(offsets[r0]);
  t.num_edges  = static_cast<int32_t>(static_cast<int64_t>(offsets[r1]) - t.edge_first);
  t.row_first  = static_cast<int32_t>(r0);
  t.num_rows   = static_cast<int32_t>(r1 - r0);
  t.aux        = 0;
  t.kind       = 0;
  t.pad        = 0;
  tiles[k]     = t;
}

// first row in [row_lo, row_hi) whose degree is < dthr (degrees are non-increasing there)
// tiles_out[dst_first[b] + k] = tiles_in[src_first[b] + k] for k < len[b]
__global__ void permute_tile_blocks_kernel(const tile_t* in, tile_t* out, const int64_t* src_first,
                                           const int64_t* dst_first, const int32_t* len)
{
  const int b = blockIdx.x;
  for (int k = threadIdx.x; k < len[b]; k += blockDim.x) out[dst_first[b] + k] = in[src_first[b] + k];
}

template <typename ET>
__global__ void degree_split_kernel(const ET* offsets, int64_t row_lo, int64_t row_hi, int64_t dthr,
                                    int64_t* out)
{
  int64_t lo = row_lo, hi = row_hi;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (static_cast<int64_t>(offsets[mid + 1]) - static_cast<int64_t>(offsets[mid]) >= dthr) lo = mid + 1;
    else hi = mid;
  }
  *out = lo;
}

// kind-0 tiles of exactly `rows_per_tile` consecutive rows (the very-low-degree tail)
template <typename ET>
__global__ void fixed_row_tiles_kernel(const ET* offsets, int64_t row_lo, int64_t row_hi,
                                       int rows_per_tile, int64_t ntiles, tile_t* tiles)
{
  int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (k >= ntiles) return;
  int64_t r0 = row_lo + k * rows_per_tile;
  int64_t r1 = min(row_hi, r0 + rows_per_tile);
  tile_t t;
  t.edge_first = static_cast<int64_t>(offsets[r0]);
  t.num_edges  = static_cast<int32_t>(static_cast<int64_t>(offsets[r1]) - t.edge_first);
  t.row_first  = static_cast<int32_t>(r0);
  t.num_rows   = static_cast<int32_t>(r1 - r0);
  t.aux        = 0;
  t.kind       = 0;
  t.pad        = 0;
  tiles[k]     = t;
}

}  // namespace

}  // namespace b200
