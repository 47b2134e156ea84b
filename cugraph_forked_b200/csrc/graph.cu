// Graph construction on the GPU: renumbering by descending major degree, degree segments,
// COO -> CSC with ascending neighbours, out-weight sums, and the tile schedule of the pull
// kernel.  B200-native replacement of
//   cpp/src/structure/renumber_edgelist_impl.cuh:96-452 (compute_renumber_map), :923-994
//   cpp/src/structure/detail/structure_utils.cuh:173-490 (compress_edgelist, sort_adjacency_list)
//   cpp/src/structure/create_graph_from_edgelist_impl.cuh:570-774 (single-GPU path)
//   cpp/src/structure/graph_view_impl.cuh:203-265 (degrees, weight sums)
// Everything is hand-written kernels + the scan / radix sort of prims.cuh.  Instead of the
// reference's atomic slot claim followed by a segmented neighbour sort, the CSC is produced
// by ONE stable radix sort of (new_major << 32 | new_minor) keys, which yields the offsets
// (from the already sorted degrees), the ascending neighbour lists and a deterministic weight
// order at once.
#include <cstdlib>

#include "banded.cuh"
#include "graph.cuh"
#include "graph_kernels.cuh"

namespace b200 {


// ------------------------------------------------------------------------------------------
// schedule
// ------------------------------------------------------------------------------------------
// The tile list is built heavy-first: chunks of the hub rows (all gather, no row work), then row
// tiles by descending degree, then runs of low-degree rows (little to gather, most row work).
// CTA c walks positions c, c + G, c + 2G, ...; taking blocks of G tiles alternately from the
// front and from the back of the list makes every CTA alternate gather-heavy and
// reduce-heavy tiles, so its gather and reduce warps are busy at the same time instead of in
// turn.  Any order is valid (chunk partials are addressed by tile.aux).
static void interleave_tiles(b200_handle* h, edge_partition& ep)
{
  const char* env = std::getenv("B200_TILE_INTERLEAVE");
  if (env && std::atoi(env) == 0) return;
  cudaStream_t s  = h->stream;
  const int64_t n = ep.num_tiles, G = h->sm_count;
  if (n < 4 * G) return;
  std::vector<int64_t> src, dst;
  std::vector<int32_t> len;
  int64_t f = 0, e = n, d = 0;
  for (int64_t b = 0; f < e; ++b) {
    const int64_t l = std::min<int64_t>(G, e - f);
    if (b % 2 == 0) { src.push_back(f); f += l; }
    else { e -= l; src.push_back(e); }
    dst.push_back(d);
    len.push_back(static_cast<int32_t>(l));
    d += l;
  }
  const size_t nb = src.size();
  dbuf<int64_t> d_src(nb, s), d_dst(nb, s);
  dbuf<int32_t> d_len(nb, s);
  CUDA_TRY(cudaMemcpyAsync(d_src.data(), src.data(), nb * 8, cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaMemcpyAsync(d_dst.data(), dst.data(), nb * 8, cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaMemcpyAsync(d_len.data(), len.data(), nb * 4, cudaMemcpyHostToDevice, s));
  dbuf<tile_t> out(static_cast<size_t>(n), s);
  permute_tile_blocks_kernel<<<static_cast<unsigned>(nb), 128, 0, s>>>(
    ep.tiles.data(), out.data(), d_src.data(), d_dst.data(), d_len.data());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  ep.tiles = std::move(out);
}

// generic variant for multi-GPU partitions, whose rows follow the GLOBAL degree order: a greedy
// pass over the offsets on the host packs complete rows into tiles of <= T edges and <= rowcap
// rows; any row with >= kLongRow edges is chunked, wherever it sits
constexpr int64_t kLongRow = kMidDegreeThreshold;

template <typename ET>
static void build_schedule_generic(b200_handle* h, edge_partition& ep, int T, int rowcap)
{
  cudaStream_t s = h->stream;
  const size_t nr = static_cast<size_t>(ep.num_rows);
  std::vector<ET> off(nr + 1, 0);
  if (nr)
    CUDA_TRY(cudaMemcpyAsync(off.data(), ep.offsets.data(), (nr + 1) * sizeof(ET),
                             cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  int64_t zero_first = 0;
  for (int64_t r = static_cast<int64_t>(nr); r > 0; --r)
    if (off[r] != off[r - 1]) { zero_first = r; break; }
  std::vector<tile_t> chunk_tiles, row_tiles;
  std::vector<int64_t> chunk_first;
  std::vector<int32_t> long_ids;
  tile_t cur{};
  auto flush = [&] {
    if (cur.num_rows > 0) row_tiles.push_back(cur);
    cur = tile_t{};
  };
  for (int64_t r = 0; r < zero_first; ++r) {
    const int64_t b = off[r], e = off[r + 1], d = e - b;
    if (d >= kLongRow) {
      flush();
      long_ids.push_back(static_cast<int32_t>(r));
      chunk_first.push_back(static_cast<int64_t>(chunk_tiles.size()));
      for (int64_t c = b; c < e; c += T) {
        tile_t t{};
        t.edge_first = c;
        t.num_edges  = static_cast<int32_t>(std::min<int64_t>(T, e - c));
        t.row_first  = static_cast<int32_t>(r);
        t.num_rows   = 1;
        t.aux        = static_cast<int32_t>(chunk_tiles.size());
        t.kind       = 1;
        chunk_tiles.push_back(t);
      }
      continue;
    }
    if (cur.num_rows == rowcap || cur.num_edges + d > T) flush();
    if (cur.num_rows == 0) {
      cur.edge_first = b;
      cur.row_first  = static_cast<int32_t>(r);
    }
    cur.num_rows += 1;
    cur.num_edges += static_cast<int32_t>(d);
  }
  flush();
  chunk_first.push_back(static_cast<int64_t>(chunk_tiles.size()));
  ep.num_chunks    = static_cast<int64_t>(chunk_tiles.size());
  ep.num_long_rows = static_cast<int64_t>(long_ids.size());
  ep.num_tiles     = ep.num_chunks + static_cast<int64_t>(row_tiles.size());
  ep.zero_first    = zero_first;
  ep.tiles.alloc(static_cast<size_t>(std::max<int64_t>(ep.num_tiles, 1)), s);
  if (!chunk_tiles.empty())
    CUDA_TRY(cudaMemcpyAsync(ep.tiles.data(), chunk_tiles.data(), chunk_tiles.size() * sizeof(tile_t),
                             cudaMemcpyHostToDevice, s));
  if (!row_tiles.empty())
    CUDA_TRY(cudaMemcpyAsync(ep.tiles.data() + chunk_tiles.size(), row_tiles.data(),
                             row_tiles.size() * sizeof(tile_t), cudaMemcpyHostToDevice, s));
  ep.long_row_chunk_first.alloc(chunk_first.size(), s);
  CUDA_TRY(cudaMemcpyAsync(ep.long_row_chunk_first.data(), chunk_first.data(),
                           chunk_first.size() * sizeof(int64_t), cudaMemcpyHostToDevice, s));
  ep.long_row_ids.alloc(std::max<size_t>(long_ids.size(), 1), s);
  if (!long_ids.empty())
    CUDA_TRY(cudaMemcpyAsync(ep.long_row_ids.data(), long_ids.data(), long_ids.size() * 4,
                             cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  interleave_tiles(h, ep);
  ep.sched_tile_edges = T;
  ep.sched_row_cap    = rowcap;
}

template <typename ET>
static void build_schedule_t(b200_handle* h, edge_partition& ep, int T, int rowcap)
{
  if (ep.generic_schedule) {
    build_schedule_generic<ET>(h, ep, T, rowcap);
    return;
  }
  cudaStream_t s = h->stream;
  const ET* off  = reinterpret_cast<const ET*>(ep.offsets.data());
  // rows [0, n_long) are split into chunks; rows [n_long, n_nz) are packed into row tiles
  int64_t n_long = ep.has_segments ? ep.seg[1] : 0;
  int64_t n_nz   = ep.has_segments ? ep.seg[3] : ep.num_rows;
  std::vector<ET> h_off(static_cast<size_t>(n_long) + 1);
  CUDA_TRY(cudaMemcpyAsync(h_off.data(), off, h_off.size() * sizeof(ET), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  std::vector<tile_t> chunk_tiles;
  std::vector<int64_t> chunk_first(static_cast<size_t>(n_long) + 1, 0);
  for (int64_t r = 0; r < n_long; ++r) {
    int64_t b = h_off[r], e = h_off[r + 1];
    chunk_first[r] = static_cast<int64_t>(chunk_tiles.size());
    for (int64_t c = b; c < e; c += T) {
      tile_t t{};
      t.edge_first = c;
      t.num_edges  = static_cast<int32_t>(std::min<int64_t>(T, e - c));
      t.row_first  = static_cast<int32_t>(r);
      t.num_rows   = 1;
      t.aux        = static_cast<int32_t>(chunk_tiles.size());
      t.kind       = 1;
      chunk_tiles.push_back(t);
    }
  }
  chunk_first[n_long] = static_cast<int64_t>(chunk_tiles.size());
  // rows [n_long, r_split) have degree >= dthr: windows of T edges hold <= rowcap rows;
  // rows [r_split, n_nz) (degree < dthr) are cut into runs of rowcap rows (< T edges)
  const int64_t dthr = div_up(T, rowcap);
  int64_t r_split    = n_nz;
  if (n_nz > n_long) {
    dbuf<int64_t> d_split(1, s);
    degree_split_kernel<ET><<<1, 1, 0, s>>>(off, n_long, n_nz, dthr, d_split.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(&r_split, d_split.data(), 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  ET h_split_off = 0;
  CUDA_TRY(cudaMemcpyAsync(&h_split_off, off + r_split, sizeof(ET), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  int64_t edge_lo     = h_off[n_long];
  int64_t edge_hi     = static_cast<int64_t>(h_split_off);
  int64_t n_row_tiles = (r_split > n_long) ? std::max<int64_t>(div_up(edge_hi - edge_lo, T), 1) : 0;
  int64_t n_fix_tiles = div_up(n_nz - r_split, rowcap);
  ep.num_chunks    = static_cast<int64_t>(chunk_tiles.size());
  ep.num_long_rows = n_long;
  ep.zero_first    = n_nz;
  ep.num_tiles     = ep.num_chunks + n_row_tiles + n_fix_tiles;
  ep.tiles.alloc(static_cast<size_t>(std::max<int64_t>(ep.num_tiles, 1)), s);
  if (!chunk_tiles.empty())
    CUDA_TRY(cudaMemcpyAsync(ep.tiles.data(), chunk_tiles.data(), chunk_tiles.size() * sizeof(tile_t),
                             cudaMemcpyHostToDevice, s));
  ep.long_row_chunk_first.alloc(chunk_first.size(), s);
  CUDA_TRY(cudaMemcpyAsync(ep.long_row_chunk_first.data(), chunk_first.data(),
                           chunk_first.size() * sizeof(int64_t), cudaMemcpyHostToDevice, s));
  if (n_row_tiles > 0) {
    row_tiles_kernel<ET><<<grid_for(n_row_tiles), kThreads, 0, s>>>(
      off, n_long, r_split, edge_lo, T, n_row_tiles, ep.tiles.data() + ep.num_chunks);
    CUDA_TRY(cudaGetLastError());
  }
  if (n_fix_tiles > 0) {
    fixed_row_tiles_kernel<ET><<<grid_for(n_fix_tiles), kThreads, 0, s>>>(
      off, r_split, n_nz, rowcap, n_fix_tiles, ep.tiles.data() + ep.num_chunks + n_row_tiles);
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  interleave_tiles(h, ep);
  ep.sched_tile_edges = T;
  ep.sched_row_cap    = rowcap;
}

void build_schedule(b200_handle* h, edge_partition& ep, int T, int rowcap)
{
  if (ep.sched_tile_edges == T && ep.sched_row_cap == rowcap) return;
  if (ep.offsets64) build_schedule_t<int64_t>(h, ep, T, rowcap);
  else build_schedule_t<int32_t>(h, ep, T, rowcap);
}

// ------------------------------------------------------------------------------------------
// schedule of a REDUCED partition (banded.cuh): rows keep their ids but their entry counts are
// no longer monotone.  Rows [0, n_long) (the original high-degree segment, a few ten thousand
// rows) are packed on the host exactly like the generic schedule; rows [n_long, n_nz) all have
// fewer than kLongRow entries and are packed on the device: window k of T entries owns the
// rows whose first entry lies in it, cut into tiles of <= rowcap rows.
// ------------------------------------------------------------------------------------------
namespace {

__global__ void window_first_rows_kernel(const int32_t* off, int64_t row_lo, int64_t row_hi,
                                         int64_t edge_lo, int T, int64_t nwin, int64_t* first_row)
{
  int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (k > nwin) return;
  if (k == nwin) { first_row[k] = row_hi; return; }
  const int64_t target = edge_lo + k * T;
  int64_t lo = row_lo, hi = row_hi;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (static_cast<int64_t>(off[mid]) < target) lo = mid + 1;
    else hi = mid;
  }
  first_row[k] = lo;
}

__global__ void window_tile_counts_kernel(const int64_t* first_row, int64_t nwin, int rowcap,
                                          uint32_t* counts)
{
  int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (k >= nwin) return;
  const int64_t rows = first_row[k + 1] - first_row[k];
  counts[k] = static_cast<uint32_t>((rows + rowcap - 1) / rowcap);
}

__global__ void window_tiles_kernel(const int32_t* off, const int64_t* first_row,
                                    const uint32_t* tile_base, int64_t nwin, int rowcap,
                                    tile_t* tiles)
{
  int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (k >= nwin) return;
  const int64_t r_end = first_row[k + 1];
  uint32_t j          = tile_base[k];
  for (int64_t r0 = first_row[k]; r0 < r_end; r0 += rowcap, ++j) {
    const int64_t r1 = min(r_end, r0 + rowcap);
    tile_t t;
    t.edge_first = static_cast<int64_t>(off[r0]);
    t.num_edges  = static_cast<int32_t>(static_cast<int64_t>(off[r1]) - t.edge_first);
    t.row_first  = static_cast<int32_t>(r0);
    t.num_rows   = static_cast<int32_t>(r1 - r0);
    t.aux        = 0;
    t.kind       = 0;
    t.pad        = 0;
    tiles[j]     = t;
  }
}

}  // namespace

void build_schedule_reduced(b200_handle* h, edge_partition& rp, int64_t n_long, int64_t n_nz, int T,
                            int rowcap)
{
  cudaStream_t s = h->stream;
  B200_EXPECTS(!rp.offsets64, kInvalidInput, "reduced partitions use 32-bit offsets");
  const int32_t* off = reinterpret_cast<const int32_t*>(rp.offsets.data());
  n_long = std::min(n_long, n_nz);
  // ---- host: rows [0, n_long) ----
  std::vector<int32_t> h_off(static_cast<size_t>(n_long) + 1, 0);
  int32_t h_end = 0;
  CUDA_TRY(cudaMemcpyAsync(h_off.data(), off, h_off.size() * 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&h_end, off + n_nz, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  std::vector<tile_t> chunk_tiles, row_tiles;
  std::vector<int64_t> chunk_first;
  std::vector<int32_t> long_ids;
  tile_t cur{};
  auto flush = [&] {
    if (cur.num_rows > 0) row_tiles.push_back(cur);
    cur = tile_t{};
  };
  for (int64_t r = 0; r < n_long; ++r) {
    const int64_t b = h_off[r], e = h_off[r + 1], d = e - b;
    if (d >= kLongRow) {
      flush();
      long_ids.push_back(static_cast<int32_t>(r));
      chunk_first.push_back(static_cast<int64_t>(chunk_tiles.size()));
      for (int64_t c = b; c < e; c += T) {
        tile_t t{};
        t.edge_first = c;
        t.num_edges  = static_cast<int32_t>(std::min<int64_t>(T, e - c));
        t.row_first  = static_cast<int32_t>(r);
        t.num_rows   = 1;
        t.aux        = static_cast<int32_t>(chunk_tiles.size());
        t.kind       = 1;
        chunk_tiles.push_back(t);
      }
      continue;
    }
    if (cur.num_rows == rowcap || cur.num_edges + d > T) flush();
    if (cur.num_rows == 0) {
      cur.edge_first = b;
      cur.row_first  = static_cast<int32_t>(r);
    }
    cur.num_rows += 1;
    cur.num_edges += static_cast<int32_t>(d);
  }
  flush();
  chunk_first.push_back(static_cast<int64_t>(chunk_tiles.size()));
  // ---- device: rows [n_long, n_nz) ----
  const int64_t edge_lo = h_off[n_long], edge_hi = h_end;
  const int64_t nwin    = (n_nz > n_long) ? std::max<int64_t>(div_up(edge_hi - edge_lo, T), 1) : 0;
  dbuf<int64_t> first_row(static_cast<size_t>(nwin) + 1, s);
  dbuf<uint32_t> counts(static_cast<size_t>(nwin) + 1, s);
  uint32_t n_win_tiles = 0;
  if (nwin > 0) {
    window_first_rows_kernel<<<grid_for(nwin + 1), kThreads, 0, s>>>(off, n_long, n_nz, edge_lo, T,
                                                                     nwin, first_row.data());
    window_tile_counts_kernel<<<grid_for(nwin), kThreads, 0, s>>>(first_row.data(), nwin, rowcap,
                                                                  counts.data());
    CUDA_TRY(cudaGetLastError());
    exclusive_scan<uint32_t, uint32_t>(counts.data(), counts.data(), static_cast<size_t>(nwin), s,
                                       true);
    CUDA_TRY(cudaMemcpyAsync(&n_win_tiles, counts.data() + nwin, 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  rp.num_chunks    = static_cast<int64_t>(chunk_tiles.size());
  rp.num_long_rows = static_cast<int64_t>(long_ids.size());
  rp.zero_first    = n_nz;
  rp.num_tiles     = rp.num_chunks + static_cast<int64_t>(row_tiles.size()) + n_win_tiles;
  rp.tiles.alloc(static_cast<size_t>(std::max<int64_t>(rp.num_tiles, 1)), s);
  if (!chunk_tiles.empty())
    CUDA_TRY(cudaMemcpyAsync(rp.tiles.data(), chunk_tiles.data(), chunk_tiles.size() * sizeof(tile_t),
                             cudaMemcpyHostToDevice, s));
  if (!row_tiles.empty())
    CUDA_TRY(cudaMemcpyAsync(rp.tiles.data() + chunk_tiles.size(), row_tiles.data(),
                             row_tiles.size() * sizeof(tile_t), cudaMemcpyHostToDevice, s));
  if (n_win_tiles > 0) {
    window_tiles_kernel<<<grid_for(nwin), kThreads, 0, s>>>(
      off, first_row.data(), counts.data(), nwin, rowcap,
      rp.tiles.data() + chunk_tiles.size() + row_tiles.size());
    CUDA_TRY(cudaGetLastError());
  }
  rp.long_row_chunk_first.alloc(chunk_first.size(), s);
  CUDA_TRY(cudaMemcpyAsync(rp.long_row_chunk_first.data(), chunk_first.data(),
                           chunk_first.size() * sizeof(int64_t), cudaMemcpyHostToDevice, s));
  rp.long_row_ids.alloc(std::max<size_t>(long_ids.size(), 1), s);
  if (!long_ids.empty())
    CUDA_TRY(cudaMemcpyAsync(rp.long_row_ids.data(), long_ids.data(), long_ids.size() * 4,
                             cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  interleave_tiles(h, rp);
  rp.sched_tile_edges = T;
  rp.sched_row_cap    = rowcap;
}

// ------------------------------------------------------------------------------------------
// single-GPU build
// ------------------------------------------------------------------------------------------
template <typename VT>
static void build_partition_from_coo(b200_handle* h, const VT* majors, const VT* minors,
                                     const void* weights, int weight_type, size_t ne,
                                     vertex_lookup<VT> lk, const uint32_t* new_of_sorted,
                                     const uint32_t* sorted_deg, int64_t nv, edge_partition& ep)
{
  // keys -> stable sort -> indices / weights; offsets = scan(sorted degrees)
  cudaStream_t s = h->stream;
  ep.num_rows    = nv;
  ep.num_edges   = static_cast<int64_t>(ne);
  ep.num_minors  = nv;
  ep.offsets64   = ne >= static_cast<size_t>(0x7fffffff);  // c_api/graph_sg.cpp: int32_threshold
  int vbits      = bits_needed(nv > 0 ? static_cast<uint64_t>(nv - 1) : 0);

  dbuf<uint64_t> k0(ne, s), k1(ne, s);
  dbuf<uint32_t> v0(ne, s), v1(ne, s);
  dbuf<int> bad(1, s);
  CUDA_TRY(cudaMemsetAsync(bad.data(), 0, sizeof(int), s));
  if (ne) {
    edge_keys_kernel<VT><<<grid_for(ne, 4), kThreads, 0, s>>>(majors, minors, ne, lk, new_of_sorted,
                                                            k0.data(), v0.data(), bad.data());
    CUDA_TRY(cudaGetLastError());
  }
  int h_bad = 0;
  CUDA_TRY(cudaMemcpyAsync(&h_bad, bad.data(), sizeof(int), cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(!h_bad, kInvalidInput,
               "Invalid input argument: edge endpoints are not in the vertex list.");
  int where = radix_sort<uint64_t>(k0.data(), k1.data(), v0.data(), v1.data(), ne,
                                   {{0, vbits}, {32, 32 + vbits}}, s);
  uint64_t* ks = where ? k1.data() : k0.data();
  uint32_t* vs = where ? v1.data() : v0.data();

  ep.indices.alloc(ne, s, 16);
  if (weights) {
    ep.weights.alloc(ne * dtype_size(weight_type), s, 32);
    if (ne) {
      if (weight_type == B200_FLOAT32)
        finish_edges_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ks, vs, ne, static_cast<const float*>(weights), ep.indices.data(),
          reinterpret_cast<float*>(ep.weights.data()));
      else
        finish_edges_kernel<double><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ks, vs, ne, static_cast<const double*>(weights), ep.indices.data(),
          reinterpret_cast<double*>(ep.weights.data()));
      CUDA_TRY(cudaGetLastError());
    }
  } else if (ne) {
    finish_edges_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(ks, vs, ne, nullptr,
                                                                   ep.indices.data(), nullptr);
    CUDA_TRY(cudaGetLastError());
  }
  // offsets
  dbuf<uint64_t> off64(static_cast<size_t>(nv) + 1, s);
  exclusive_scan<uint32_t, uint64_t>(sorted_deg, off64.data(), static_cast<size_t>(nv), s, true);
  if (ep.offsets64) {
    ep.offsets.alloc((static_cast<size_t>(nv) + 1) * 8, s, 64);
    CUDA_TRY(cudaMemcpyAsync(ep.offsets.data(), off64.data(), (static_cast<size_t>(nv) + 1) * 8,
                             cudaMemcpyDeviceToDevice, s));
  } else {
    ep.offsets.alloc((static_cast<size_t>(nv) + 1) * 4, s, 64);
    narrow_offsets_kernel<int32_t><<<grid_for(nv + 1, 4), kThreads, 0, s>>>(
      off64.data(), static_cast<size_t>(nv) + 1, reinterpret_cast<int32_t*>(ep.offsets.data()));
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
}

template <typename VT>
static void create_sg_t(b200_handle* h, const VT* src, const VT* dst, const void* weights,
                        size_t ne, int weight_type, const VT* vertices, size_t n_given,
                        bool store_transposed, bool renumber, b200_graph* g)
{
  cudaStream_t s   = h->stream;
  const VT* majors = store_transposed ? dst : src;
  const VT* minors = store_transposed ? src : dst;

  // ---- 1. sorted unique vertex list (renumber_edgelist_impl.cuh:96-217) ----
  dbuf<unsigned long long> d_max(1, s);
  dbuf<int> d_flag(1, s);
  CUDA_TRY(cudaMemsetAsync(d_max.data(), 0, 8, s));
  CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
  if (ne) {
    max_id_kernel<VT><<<std::min(grid_for(ne, 8), 4096u), kThreads, 0, s>>>(src, dst, ne, d_max.data(),
                                                                        d_flag.data());
    CUDA_TRY(cudaGetLastError());
  }
  if (vertices && n_given) {
    max_id_kernel<VT><<<std::min(grid_for(n_given, 8), 4096u), kThreads, 0, s>>>(
      vertices, nullptr, n_given, d_max.data(), d_flag.data());
    CUDA_TRY(cudaGetLastError());
  }
  unsigned long long h_max = 0;
  int h_flag               = 0;
  CUDA_TRY(cudaMemcpyAsync(&h_max, d_max.data(), 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(!h_flag, kInvalidInput, "Invalid input argument: vertex ids should be non-negative.");
  bool empty = (ne == 0) && !(vertices && n_given);

  int64_t nv = 0;
  dbuf<VT> sorted_vertices;
  dbuf<uint32_t> pos;  // direct map id -> position (mode 1)
  vertex_lookup<VT> lk{0, nullptr, nullptr, 0};
  if (empty) {
    nv = 0;
  } else if (!renumber) {
    // ids are the internal ids; V = max id + 1 (or the given list, which must be 0..V-1)
    nv = static_cast<int64_t>(h_max) + 1;
    if (vertices && n_given) nv = std::max<int64_t>(nv, static_cast<int64_t>(n_given));
    lk = {0, nullptr, nullptr, nv};
  } else {
    size_t span = static_cast<size_t>(h_max) + 1;
    bool identity = false;
    if (vertices && n_given == span) {
      CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
      is_identity_kernel<VT><<<std::min(grid_for(n_given, 8), 4096u), kThreads, 0, s>>>(vertices, n_given,
                                                                                    d_flag.data());
      CUDA_TRY(cudaGetLastError());
      CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
      CUDA_TRY(cudaStreamSynchronize(s));
      identity = !h_flag;
    }
    if (identity) {
      nv = static_cast<int64_t>(span);
      lk = {0, nullptr, nullptr, nv};
    } else if (span <= std::max<size_t>(size_t{1} << 26, 8 * (ne + n_given))) {
      // dense-ish id space: presence flags + scan (same result as sort + unique)
      dbuf<uint32_t> flags(span, s);
      CUDA_TRY(cudaMemsetAsync(flags.data(), 0, span * 4, s));
      if (vertices && n_given) {
        mark_present_kernel<VT><<<grid_for(n_given, 4), kThreads, 0, s>>>(vertices, nullptr, n_given,
                                                                        flags.data());
      } else {
        mark_present_kernel<VT><<<grid_for(ne, 4), kThreads, 0, s>>>(src, dst, ne, flags.data());
      }
      CUDA_TRY(cudaGetLastError());
      pos.alloc(span + 1, s);
      exclusive_scan<uint32_t, uint32_t>(flags.data(), pos.data(), span, s, true);
      uint32_t h_nv = 0;
      CUDA_TRY(cudaMemcpyAsync(&h_nv, pos.data() + span, 4, cudaMemcpyDeviceToHost, s));
      CUDA_TRY(cudaStreamSynchronize(s));
      nv = h_nv;
      sorted_vertices.alloc(static_cast<size_t>(nv), s);
      compact_present_kernel<VT><<<grid_for(span, 4), kThreads, 0, s>>>(flags.data(), pos.data(), span,
                                                                      sorted_vertices.data());
      CUDA_TRY(cudaGetLastError());
      CUDA_TRY(cudaStreamSynchronize(s));
      lk = {1, pos.data(), sorted_vertices.data(), nv};
    } else {
      // sparse id space: radix sort + unique, lookups by binary search
      size_t n_all = (vertices && n_given) ? n_given : 2 * ne;
      dbuf<uint64_t> a(n_all, s), b(n_all, s);
      {
        dbuf<VT> tmp;  // widen to u64 keys
        if (vertices && n_given) launch_widen_ids(vertices, nullptr, n_given, sizeof(VT), a.data(), s);
        else launch_widen_ids(src, dst, ne, sizeof(VT), a.data(), s);
      }
      int where = radix_sort<uint64_t>(a.data(), b.data(), nullptr, nullptr, n_all,
                                       {{0, bits_needed(h_max)}}, s);
      uint64_t* srt = where ? b.data() : a.data();
      sorted_vertices.alloc(n_all, s);
      nv = launch_unique_ids(srt, n_all, sizeof(VT), sorted_vertices.data(), s);
      lk = {2, nullptr, sorted_vertices.data(), nv};
    }
  }
  B200_EXPECTS(nv < (int64_t{1} << 31), kInvalidInput,
               "number of vertices must fit 32-bit internal ids");

  g->num_vertices = nv;
  g->num_edges    = static_cast<int64_t>(ne);
  g->local_first  = 0;
  g->local_last   = nv;
  g->renumbered   = renumber;
  g->parts.resize(1);
  edge_partition& ep = g->parts[0];

  // ---- 2. major degrees (renumber_edgelist_impl.cuh:219-382) ----
  dbuf<uint32_t> deg(static_cast<size_t>(nv), s);
  CUDA_TRY(cudaMemsetAsync(deg.data(), 0, static_cast<size_t>(nv) * 4 + (nv == 0 ? 4 : 0), s));
  CUDA_TRY(cudaMemsetAsync(d_flag.data(), 0, 4, s));
  if (ne) {
    count_degree_kernel<VT><<<grid_for(ne, 4), kThreads, 0, s>>>(majors, ne, lk, deg.data(),
                                                               d_flag.data());
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaMemcpyAsync(&h_flag, d_flag.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(!h_flag, kInvalidInput,
               "Invalid input argument: edge endpoints are not in the vertex list.");

  dbuf<uint32_t> new_of_sorted;
  dbuf<uint32_t> sorted_deg(static_cast<size_t>(nv), s);
  g->renumber_map.alloc(static_cast<size_t>(nv) * sizeof(VT), s);
  VT* labels = reinterpret_cast<VT*>(g->renumber_map.data());

  if (renumber && nv > 0) {
    // ---- 3. stable sort by descending degree (renumber_edgelist_impl.cuh:384-390) ----
    dbuf<uint32_t> d_maxdeg(1, s);
    CUDA_TRY(cudaMemsetAsync(d_maxdeg.data(), 0, 4, s));
    max_u32_kernel<<<std::min(grid_for(nv, 8), 4096u), kThreads, 0, s>>>(deg.data(), nv, d_maxdeg.data());
    CUDA_TRY(cudaGetLastError());
    uint32_t maxdeg = 0;
    CUDA_TRY(cudaMemcpyAsync(&maxdeg, d_maxdeg.data(), 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    dbuf<uint32_t> k0(nv, s), k1(nv, s), v0(nv, s), v1(nv, s);
    degree_keys_kernel<<<grid_for(nv, 4), kThreads, 0, s>>>(deg.data(), nv, maxdeg, k0.data(), v0.data());
    CUDA_TRY(cudaGetLastError());
    int where = radix_sort<uint32_t>(k0.data(), k1.data(), v0.data(), v1.data(), nv,
                                     {{0, bits_needed(maxdeg)}}, s);
    new_of_sorted.alloc(nv, s);
    finish_renumber_kernel<VT><<<grid_for(nv, 4), kThreads, 0, s>>>(
      where ? k1.data() : k0.data(), where ? v1.data() : v0.data(), nv, maxdeg,
      sorted_vertices.data(), lk.mode == 0, labels, new_of_sorted.data(), sorted_deg.data());
    CUDA_TRY(cudaGetLastError());

    // ---- 4. segment offsets (renumber_edgelist_impl.cuh:392-449) ----
    std::vector<int64_t> thr{kMidDegreeThreshold, kLowDegreeThreshold, 1};
    dbuf<int64_t> d_thr(thr.size(), s), d_seg(thr.size(), s);
    CUDA_TRY(cudaMemcpyAsync(d_thr.data(), thr.data(), thr.size() * 8, cudaMemcpyHostToDevice, s));
    segment_bounds_kernel<<<1, 32, 0, s>>>(sorted_deg.data(), nv, d_thr.data(),
                                           static_cast<int>(thr.size()), d_seg.data());
    CUDA_TRY(cudaGetLastError());
    std::vector<int64_t> h_seg(thr.size());
    CUDA_TRY(cudaMemcpyAsync(h_seg.data(), d_seg.data(), thr.size() * 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    g->num_segments       = 4;
    g->segment_offsets[0] = 0;
    for (int i = 0; i < 3; ++i) g->segment_offsets[i + 1] = h_seg[i];
    g->segment_offsets[4] = nv;
    ep.has_segments       = true;
    for (int i = 0; i < 5; ++i) ep.seg[i] = g->segment_offsets[i];
  } else {
    if (nv > 0) {
      iota_kernel<VT><<<grid_for(nv, 4), kThreads, 0, s>>>(labels, nv, VT(0));
      CUDA_TRY(cudaGetLastError());
      CUDA_TRY(cudaMemcpyAsync(sorted_deg.data(), deg.data(), static_cast<size_t>(nv) * 4,
                               cudaMemcpyDeviceToDevice, s));
    }
    g->num_segments = 0;
    ep.has_segments = false;
  }

  // ---- 5. renumber edges + compress + neighbour sort ----
  build_partition_from_coo<VT>(h, majors, minors, weights, weight_type, ne, lk,
                               (renumber && nv > 0) ? new_of_sorted.data() : nullptr,
                               sorted_deg.data(), nv, ep);
  ep.major_first = 0;
  ep.minor_first = 0;
  if (!ep.has_segments) { ep.seg[0] = 0; ep.seg[1] = 0; ep.seg[2] = 0; ep.seg[3] = nv; ep.seg[4] = nv; }
  g->edge_type = ep.offsets64 ? B200_INT64 : B200_INT32;
}

void graph_out_weight_sums(b200_handle* h, b200_graph* g)
{
  if (g->have_out_weight_sums) return;
  cudaStream_t s = h->stream;
  if (g->multi_gpu) {
    graph_out_weight_sums_mg(h, g);
    return;
  }
  edge_partition& ep = g->parts[0];
  size_t nv          = static_cast<size_t>(g->num_vertices);
  size_t ne          = static_cast<size_t>(ep.num_edges);
  g->out_weight_sums.alloc(nv, s);
  if (g->weighted) {
    CUDA_TRY(cudaMemsetAsync(g->out_weight_sums.data(), 0, std::max<size_t>(nv, 1) * 8, s));
    if (ne) {
      if (g->weight_type == B200_FLOAT32)
        out_weight_kernel<float><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ep.indices.data(), reinterpret_cast<const float*>(ep.weights.data()), ne,
          g->out_weight_sums.data());
      else
        out_weight_kernel<double><<<grid_for(ne, 4), kThreads, 0, s>>>(
          ep.indices.data(), reinterpret_cast<const double*>(ep.weights.data()), ne,
          g->out_weight_sums.data());
      CUDA_TRY(cudaGetLastError());
    }
  } else {
    dbuf<unsigned long long> cnt(nv, s);
    CUDA_TRY(cudaMemsetAsync(cnt.data(), 0, std::max<size_t>(nv, 1) * 8, s));
    if (ne) {
      out_degree_kernel<<<grid_for(ne, 4), kThreads, 0, s>>>(ep.indices.data(), ne, cnt.data());
      CUDA_TRY(cudaGetLastError());
    }
    if (nv) {
      u64_to_double_kernel<<<grid_for(nv, 4), kThreads, 0, s>>>(cnt.data(), nv,
                                                                g->out_weight_sums.data());
      CUDA_TRY(cudaGetLastError());
    }
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  g->have_out_weight_sums = true;
}

// small helpers referenced above (kept out of templates to have one definition)
__global__ void widen_ids_kernel32(const int32_t* a, const int32_t* b, size_t n, uint64_t* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    out[i] = static_cast<uint64_t>(a[i]);
    if (b) out[n + i] = static_cast<uint64_t>(b[i]);
  }
}
__global__ void widen_ids_kernel64(const int64_t* a, const int64_t* b, size_t n, uint64_t* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    out[i] = static_cast<uint64_t>(a[i]);
    if (b) out[n + i] = static_cast<uint64_t>(b[i]);
  }
}
void launch_widen_ids(const void* a, const void* b, size_t n, int width, uint64_t* out,
                      cudaStream_t s)
{
  if (n == 0) return;
  if (width == 4)
    widen_ids_kernel32<<<grid_for(n, 4), kThreads, 0, s>>>(static_cast<const int32_t*>(a),
                                                          static_cast<const int32_t*>(b), n, out);
  else
    widen_ids_kernel64<<<grid_for(n, 4), kThreads, 0, s>>>(static_cast<const int64_t*>(a),
                                                          static_cast<const int64_t*>(b), n, out);
  CUDA_TRY(cudaGetLastError());
}
__global__ void unique_flags_kernel(const uint64_t* sorted, size_t n, uint32_t* flags)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    flags[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1u : 0u;
}
template <typename VT>
__global__ void unique_compact_kernel(const uint64_t* sorted, const uint32_t* flags,
                                      const uint32_t* pos, size_t n, VT* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    if (flags[i]) out[pos[i]] = static_cast<VT>(sorted[i]);
}
int64_t launch_unique_ids(const uint64_t* sorted, size_t n, int width, void* out, cudaStream_t s)
{
  if (n == 0) return 0;
  dbuf<uint32_t> flags(n, s), pos(n + 1, s);
  unique_flags_kernel<<<grid_for(n, 4), kThreads, 0, s>>>(sorted, n, flags.data());
  CUDA_TRY(cudaGetLastError());
  exclusive_scan<uint32_t, uint32_t>(flags.data(), pos.data(), n, s, true);
  uint32_t cnt = 0;
  CUDA_TRY(cudaMemcpyAsync(&cnt, pos.data() + n, 4, cudaMemcpyDeviceToHost, s));
  if (width == 4)
    unique_compact_kernel<int32_t><<<grid_for(n, 4), kThreads, 0, s>>>(sorted, flags.data(), pos.data(),
                                                                     n, static_cast<int32_t*>(out));
  else
    unique_compact_kernel<int64_t><<<grid_for(n, 4), kThreads, 0, s>>>(sorted, flags.data(), pos.data(),
                                                                     n, static_cast<int64_t*>(out));
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  return cnt;
}

void graph_create_sg(b200_handle* h, const void* src, const void* dst, const void* weights,
                     size_t ne, int vertex_type, int weight_type, const void* vertices,
                     size_t n_given, bool store_transposed, bool renumber, b200_graph* g)
{
  g->vertex_type      = vertex_type;
  g->weight_type      = weights ? weight_type : B200_FLOAT32;
  g->weighted         = weights != nullptr;
  g->store_transposed = store_transposed;
  g->multi_gpu        = false;
  if (vertex_type == B200_INT32)
    create_sg_t<int32_t>(h, static_cast<const int32_t*>(src), static_cast<const int32_t*>(dst),
                         weights, ne, weight_type, static_cast<const int32_t*>(vertices), n_given,
                         store_transposed, renumber, g);
  else
    create_sg_t<int64_t>(h, static_cast<const int64_t*>(src), static_cast<const int64_t*>(dst),
                         weights, ne, weight_type, static_cast<const int64_t*>(vertices), n_given,
                         store_transposed, renumber, g);
}

void graph_in_degrees(b200_handle* h, const b200_graph* g, void* out)
{
  const edge_partition& ep = g->parts[0];
  size_t nv                = static_cast<size_t>(ep.num_rows);
  if (nv == 0) return;
  if (ep.offsets64)
    in_degree_kernel<int64_t><<<grid_for(nv, 4), kThreads, 0, h->stream>>>(
      reinterpret_cast<const int64_t*>(ep.offsets.data()), nv, static_cast<int64_t*>(out));
  else
    in_degree_kernel<int32_t><<<grid_for(nv, 4), kThreads, 0, h->stream>>>(
      reinterpret_cast<const int32_t*>(ep.offsets.data()), nv, static_cast<int32_t*>(out));
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(h->stream));
}


// ------------------------------------------------------------------------------------------
// external id -> internal id (renumber_local_ext_vertices / collect_local_vertex_values_from_
// ext_vertex_value_pairs of the reference's C API glue, c_api/pagerank.cpp:125-210)
// ------------------------------------------------------------------------------------------
namespace {
template <typename VT>
__global__ void map_keys_kernel(const VT* labels, size_t n, uint64_t* keys, uint32_t* vals)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    keys[i] = static_cast<uint64_t>(labels[i]);
    vals[i] = static_cast<uint32_t>(i);
  }
}
template <typename VT>
__global__ void lookup_kernel(const uint64_t* sorted_ext, const uint32_t* sorted_int, size_t n_map,
                              const VT* query, size_t n, int32_t* out)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    uint64_t q = static_cast<uint64_t>(query[i]);
    size_t lo = 0, hi = n_map;
    while (lo < hi) {
      size_t mid = (lo + hi) >> 1;
      if (sorted_ext[mid] < q) lo = mid + 1;
      else hi = mid;
    }
    out[i] = (lo < n_map && sorted_ext[lo] == q && query[i] >= 0) ? static_cast<int32_t>(sorted_int[lo])
                                                                   : -1;
  }
}
template <typename VT, typename WT>
__global__ void decompress_kernel(const void* offsets_v, int offsets64, const int32_t* indices,
                                  const WT* w_in, const VT* labels, int64_t num_rows,
                                  int64_t minor_first, int majors_are_dst, VT* src, VT* dst,
                                  WT* w_out)
{
  // one warp per row
  const int lane     = threadIdx.x & 31;
  const int64_t warp = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nw   = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < num_rows; r += nw) {
    int64_t b, e;
    if (offsets64) {
      b = static_cast<const int64_t*>(offsets_v)[r];
      e = static_cast<const int64_t*>(offsets_v)[r + 1];
    } else {
      b = static_cast<const int32_t*>(offsets_v)[r];
      e = static_cast<const int32_t*>(offsets_v)[r + 1];
    }
    VT major = labels[r];
    for (int64_t i = b + lane; i < e; i += 32) {
      VT minor = labels[minor_first + indices[i]];
      src[i]   = majors_are_dst ? minor : major;
      dst[i]   = majors_are_dst ? major : minor;
      if (w_in) w_out[i] = w_in[i];
    }
  }
}
}  // namespace

void graph_ext_to_internal(b200_handle* h, b200_graph* g, const void* ext_ids, size_t n,
                           int32_t* out)
{
  cudaStream_t s = h->stream;
  size_t nloc    = static_cast<size_t>(g->local_last - g->local_first);
  if (!g->have_ext_map) {
    dbuf<uint64_t> k0(nloc, s), k1(nloc, s);
    dbuf<uint32_t> v0(nloc, s), v1(nloc, s);
    if (nloc) {
      if (g->vertex_type == B200_INT32)
        map_keys_kernel<int32_t><<<grid_for(nloc, 4), kThreads, 0, s>>>(
          reinterpret_cast<const int32_t*>(g->renumber_map.data()), nloc, k0.data(), v0.data());
      else
        map_keys_kernel<int64_t><<<grid_for(nloc, 4), kThreads, 0, s>>>(
          reinterpret_cast<const int64_t*>(g->renumber_map.data()), nloc, k0.data(), v0.data());
      CUDA_TRY(cudaGetLastError());
    }
    int where = radix_sort<uint64_t>(k0.data(), k1.data(), v0.data(), v1.data(), nloc, {{0, 63}}, s);
    g->ext_sorted   = where ? std::move(k1) : std::move(k0);
    g->ext_internal = where ? std::move(v1) : std::move(v0);
    g->have_ext_map = true;
  }
  if (n == 0) return;
  if (g->vertex_type == B200_INT32)
    lookup_kernel<int32_t><<<grid_for(n, 2), kThreads, 0, s>>>(
      g->ext_sorted.data(), g->ext_internal.data(), nloc, static_cast<const int32_t*>(ext_ids), n, out);
  else
    lookup_kernel<int64_t><<<grid_for(n, 2), kThreads, 0, s>>>(
      g->ext_sorted.data(), g->ext_internal.data(), nloc, static_cast<const int64_t*>(ext_ids), n, out);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
}

// Rebuilds a single-GPU graph with the other storage orientation, in place (what
// cugraph::c_api::transpose_storage does before a pull algorithm, c_api/graph.hpp:44-79):
// decompress to an external-id COO, then run the normal build with the full vertex list.
void graph_transpose_storage(b200_handle* h, b200_graph* g)
{
  B200_EXPECTS(!g->multi_gpu, kNotImplemented, "transpose_storage of a multi-GPU graph");
  cudaStream_t s     = h->stream;
  edge_partition& ep = g->parts[0];
  size_t ne          = static_cast<size_t>(ep.num_edges);
  size_t nv          = static_cast<size_t>(g->num_vertices);
  int vs             = dtype_size(g->vertex_type);
  int ws             = dtype_size(g->weight_type);
  dbuf<char> src(ne * vs, s), dst(ne * vs, s), w(g->weighted ? ne * ws : 0, s);
  if (ne) {
    unsigned grid = grid_for(nv * 32, 1);
    if (grid > 65535u * 8) grid = 65535u * 8;
    auto run = [&](auto vt, auto wt) {
      using VT = decltype(vt);
      using WT = decltype(wt);
      decompress_kernel<VT, WT><<<grid, kThreads, 0, s>>>(
        ep.offsets.data(), ep.offsets64 ? 1 : 0, ep.indices.data(),
        g->weighted ? reinterpret_cast<const WT*>(ep.weights.data()) : nullptr,
        reinterpret_cast<const VT*>(g->renumber_map.data()), ep.num_rows, ep.minor_first,
        g->store_transposed ? 1 : 0, reinterpret_cast<VT*>(src.data()),
        reinterpret_cast<VT*>(dst.data()), reinterpret_cast<WT*>(w.data()));
    };
    if (g->vertex_type == B200_INT32) {
      if (g->weight_type == B200_FLOAT32) run(int32_t{}, float{}); else run(int32_t{}, double{});
    } else {
      if (g->weight_type == B200_FLOAT32) run(int64_t{}, float{}); else run(int64_t{}, double{});
    }
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  b200_graph fresh;
  // the vertex list keeps isolated vertices; ids stay as they are when the graph was not renumbered
  graph_create_sg(h, src.data(), dst.data(), g->weighted ? w.data() : nullptr, ne, g->vertex_type,
                  g->weight_type, g->renumber_map.data(), nv, !g->store_transposed, g->renumbered,
                  &fresh);
  *g = std::move(fresh);
}

}  // namespace b200
