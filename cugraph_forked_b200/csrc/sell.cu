// Build of the round-2 "lane owns a segment" layout (sell.cuh) from a CSC rectangle, on the GPU
// with the hand-written scan / radix sort of prims.cuh.  32-bit offsets.
//
//   1. cell starts: edge e starts a cell if it is the first edge of its row, or its source lies
//      in another band than the source of edge e - 1, or e is a multiple of kSellCellMax;
//   2. a cell is a CANDIDATE if it has >= kSellMinCell edges; a band is HOT if its candidate cells
//      hold >= kHotBandMin edges; the cells of hot bands are pre-reduced;
//   3. two scans give every edge its position in the reduced CSC (kept: all direct edges + the
//      first edge of every pre-reduced cell) and every pre-reduced cell its partial slot
//      (row-major, so a row's partials are consecutive); the reduced CSC is written in one pass;
//   4. phase 1: one record per cell (band, length, first edge); the CELLS (not the edges: 0.06
//      cells per edge) are sorted by (band, length descending) with the stable radix sort, packed
//      32 to a slice, and the slices' interleaved 16-bit index arrays are written by one warp
//      per slice straight from the CSC (a cell's edges are consecutive there);
//   5. phase 2: the rows of the reduced CSC are counting-sorted by length inside windows of
//      kSellWindow rows (one CTA per window, stable), packed 32 to a slice and written
//      interleaved; rows with more than kSellLong entries become chunks in slices of their own;
//   6. work lists: phase-1 tasks (runs of slices of one band, ~kSellQ1 load groups) and units
//      (tasks of one band for one CTA, guided self-scheduling); phase-2 tasks (runs of slices).
#include <algorithm>
#include <chrono>
#include <cstdlib>

#include "graph.cuh"
#include "sell.cuh"

namespace b200 {

namespace {

constexpr int kBT = 256;
inline unsigned sgrid(size_t n, size_t cap = 148 * 32)
{
  size_t b = (n + kBT - 1) / kBT;
  if (b > cap) b = cap;
  if (b == 0) b = 1;
  return static_cast<unsigned>(b);
}

struct s8_keep { uint8_t v; __host__ __device__ operator uint32_t() const { return v & 1u; } };
struct s8_head { uint8_t v; __host__ __device__ operator uint32_t() const { return (v >> 1) & 1u; } };

__global__ void sell_row_starts_kernel(const int32_t* off, int64_t n_rows, uint8_t* f8)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < n_rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int32_t b = off[r], e = off[r + 1];
    if (b < e) f8[b] = 1;
  }
}

// f8[e]: bit 0 row start (input), bit 1 cell start (output)
__global__ void sell_cell_starts_kernel(const int32_t* idx, int64_t ne, int band_entries, int cell_max,
                                        uint8_t* f8)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const uint8_t rs = f8[e] & 1u;
    bool f = rs || e == 0 || (e % cell_max) == 0;
    if (!f) f = (idx[e] / band_entries) != (idx[e - 1] / band_entries);
    f8[e] = static_cast<uint8_t>(rs | (f ? 2u : 0u));
  }
}

// c8[e] = 1 if the cell of edge e has >= kSellMinCell edges; hist[band] += 1 for those edges
constexpr int kSellMaxBands = 8192;

__global__ void sell_classify_kernel(const uint8_t* f8, const int32_t* idx, int64_t ne,
                                     int band_entries, int nb, uint8_t* c8, unsigned long long* hist)
{
  extern __shared__ uint32_t s_hist[];
  for (int k = threadIdx.x; k < nb; k += blockDim.x) s_hist[k] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  const int64_t rounds = (ne + stride - 1) / stride;
  for (int64_t k = 0; k < rounds; ++k) {
    const int64_t e = k * stride + blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
    bool cand = false;
    int band  = -1;
    if (e < ne) {
      band = idx[e] / band_entries;
      int64_t p = -1;   // nearest cell start at or before e, within kSellMinCell - 1 edges
#pragma unroll
      for (int j = 0; j < kSellMinCell; ++j)
        if (p < 0 && e - j >= 0 && (f8[e - j] & 2u)) p = e - j;
      cand = true;      // no start found: the cell started >= kSellMinCell edges ago
      if (p >= 0) {
#pragma unroll
        for (int j = 1; j < kSellMinCell; ++j) {
          const int64_t q = p + j;
          if (q >= ne || (f8[q] & 2u)) cand = false;
        }
      }
      c8[e] = cand ? 1 : 0;
    }
    const int prev       = __shfl_up_sync(0xffffffffu, band, 1);
    const bool head      = lane == 0 || prev != band;
    const uint32_t heads = __ballot_sync(0xffffffffu, head);
    const uint32_t cands = __ballot_sync(0xffffffffu, cand);
    if (head && band >= 0) {
      const uint32_t above = heads & ~((2u << lane) - 1u);
      const int end        = above ? __ffs(above) - 1 : 32;
      const uint32_t run   = (end >= 32 ? 0xffffffffu : ((1u << end) - 1u)) & ~((1u << lane) - 1u);
      const int n          = __popc(cands & run);
      if (n) atomicAdd(&s_hist[band], static_cast<uint32_t>(n));
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nb; k += blockDim.x)
    if (s_hist[k]) atomicAdd(hist + k, static_cast<unsigned long long>(s_hist[k]));
}

__global__ void sell_mark_hot_kernel(const unsigned long long* hist, int nb, long long min_edges,
                                     uint8_t* hot8)
{
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nb; k += gridDim.x * blockDim.x)
    hot8[k] = hist[k] >= static_cast<unsigned long long>(min_edges) ? 1 : 0;
}

// m8[e]: bit 0 keep (entry of the reduced CSC), bit 1 head (first edge of a pre-reduced cell),
// bit 2 pre-reduced, bit 3 cell start
__global__ void sell_masks_kernel(const uint8_t* f8, const uint8_t* c8, const uint8_t* hot8,
                                  const int32_t* idx, int64_t ne, int band_entries, uint8_t* m8)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const bool f      = (f8[e] & 2u) != 0;
    const bool banded = c8[e] && hot8[idx[e] / band_entries];
    const bool keep   = !banded || f;
    const bool head   = banded && f;
    m8[e] = static_cast<uint8_t>((keep ? 1u : 0u) | (head ? 2u : 0u) | (banded ? 4u : 0u) |
                                 (f ? 8u : 0u));
  }
}

template <typename RT>
__global__ void sell_fill_reduced_kernel(const uint8_t* m8, const uint32_t* newpos,
                                         const uint32_t* pid, const int32_t* idx, const RT* w,
                                         int64_t ne, int32_t x_split, int32_t* ridx, RT* rw)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const uint8_t m = m8[e];
    if (m & 1u) {
      const bool banded = (m & 4u) != 0;
      const uint32_t j  = newpos[e];
      ridx[j] = banded ? static_cast<int32_t>(x_split + static_cast<int32_t>(pid[e])) : idx[e];
      if (w) rw[j] = banded ? RT(1) : w[e];
    }
  }
}

__global__ void sell_reduced_offsets_kernel(const int32_t* off, int64_t n, const uint32_t* newpos,
                                            int32_t* roff)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < n;
       r += int64_t(gridDim.x) * blockDim.x)
    roff[r] = static_cast<int32_t>(newpos[off[r]]);
}

// one record per pre-reduced cell, at its partial slot q (row-major): first edge, and the sort
// key (band << 32 | kSellCellMax - length); total[0] += edges of pre-reduced cells
__global__ void sell_cell_records_kernel(const uint8_t* m8, const uint32_t* pid, const int32_t* idx,
                                         int64_t ne, int band_entries, int cell_max, uint32_t* ce0,
                                         uint64_t* keys,
                                         uint32_t* vals, unsigned long long* total)
{
  unsigned long long mine = 0;
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    if (!(m8[e] & 2u)) continue;
    int len = 1;
    while (e + len < ne && !(m8[e + len] & 8u)) ++len;
    const uint32_t q = pid[e];
    ce0[q]  = static_cast<uint32_t>(e);
    keys[q] = (static_cast<uint64_t>(static_cast<uint32_t>(idx[e] / band_entries)) << 32) |
              static_cast<uint64_t>(cell_max - len);
    vals[q] = q;
    mine += static_cast<unsigned long long>(len);
  }
  for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(total, mine);
}

// bstart[k] = first sorted cell whose band is >= k, k in [0, nb]
__global__ void sell_band_starts_kernel(const uint64_t* keys, int64_t n, int nb, int64_t* bstart)
{
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= nb; k += gridDim.x * blockDim.x) {
    const uint64_t target = static_cast<uint64_t>(k) << 32;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < target) lo = mid + 1; else hi = mid;
    }
    bstart[k] = lo;
  }
}

// band of slice s: largest k with slice_first[k] <= s
__device__ __forceinline__ int sell_band_of_slice(const int64_t* slice_first, int nb, int64_t s)
{
  int lo = 0, hi = nb;   // slice_first[lo] <= s < slice_first[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (slice_first[mid] <= s) lo = mid; else hi = mid;
  }
  return lo;
}

// load groups of every phase-1 slice = ceil(length of its first (longest) cell / kSellG)
__global__ void sell_p1_groups_kernel(const uint64_t* keys, const int64_t* bstart,
                                      const int64_t* slice_first, int nb, int64_t nslices,
                                      int cell_max, uint32_t* groups)
{
  for (int64_t s = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; s < nslices;
       s += int64_t(gridDim.x) * blockDim.x) {
    const int k      = sell_band_of_slice(slice_first, nb, s);
    const int64_t c0 = bstart[k] + (s - slice_first[k]) * 32;
    const int len    = cell_max - static_cast<int>(keys[c0] & 0xffffffffu);
    groups[s]        = static_cast<uint32_t>((len + kSellG - 1) / kSellG);
  }
}

// one warp per slice: lane l copies cell l's band-local indices (and weights) into the
// interleaved arrays, kSellG at a time, padding with the band's zero entry
template <typename RT>
__global__ void sell_p1_fill_kernel(const uint64_t* keys, const uint32_t* vals, const uint32_t* ce0,
                                    const int64_t* bstart, const int64_t* slice_first, int nb,
                                    int64_t nslices, const uint32_t* sbase, const int32_t* idx,
                                    const RT* w, int band_entries, int cell_max, uint16_t* out_idx,
                                    RT* out_w, int32_t* out_slot, uint32_t* inv)
{
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarp = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t s = warp0; s < nslices; s += nwarp) {
    const int k      = sell_band_of_slice(slice_first, nb, s);
    const int64_t c  = bstart[k] + (s - slice_first[k]) * 32 + lane;
    const bool live  = c < bstart[k + 1];
    int len = 0;
    int64_t e0 = 0;
    int32_t slot = -1;
    if (live) {
      len  = cell_max - static_cast<int>(keys[c] & 0xffffffffu);
      slot = static_cast<int32_t>(vals[c]);
      e0   = ce0[slot];
      if (inv) {   // band-major partial slots: the cell's rank in the sorted order
        inv[slot] = static_cast<uint32_t>(c);
        slot      = static_cast<int32_t>(c);
      }
    }
    out_slot[s * 32 + lane] = slot;
    const uint32_t g0 = sbase[s], g1 = sbase[s + 1];
    const int32_t base = k * band_entries;
    for (uint32_t g = g0; g < g1; ++g) {
      const int j0 = static_cast<int>(g - g0) * kSellG;
      uint32_t v[kSellG];
#pragma unroll
      for (int t = 0; t < kSellG; ++t)
        v[t] = (j0 + t < len) ? static_cast<uint32_t>(idx[e0 + j0 + t] - base)
                              : static_cast<uint32_t>(band_entries);
      uint2 pk;
      pk.x = v[0] | (v[1] << 16);
      pk.y = v[2] | (v[3] << 16);
      reinterpret_cast<uint2*>(out_idx)[static_cast<int64_t>(g) * 32 + lane] = pk;
      if (w) {
#pragma unroll
        for (int t = 0; t < kSellG; ++t)
          out_w[(static_cast<int64_t>(g) * 32 + lane) * kSellG + t] =
            (j0 + t < len) ? w[e0 + j0 + t] : RT(0);
      }
    }
  }
}

// phase-1 task t of band k covers the slices whose first load group lies in
// [gfirst + t * kSellQ1, gfirst + (t + 1) * kSellQ1), gfirst = first group of the band
__global__ void sell_p1_tasks_kernel(const int64_t* task_first_of_band, const int64_t* slice_first,
                                     int nb, const uint32_t* sbase, int64_t ntasks, int64_t quantum,
                                     sell_task* tasks)
{
  for (int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; t < ntasks;
       t += int64_t(gridDim.x) * blockDim.x) {
    const int k      = sell_band_of_slice(task_first_of_band, nb, t);
    const int64_t s0 = slice_first[k], s1 = slice_first[k + 1];
    const uint32_t gfirst = sbase[s0];
    auto first_slice = [&](int64_t tt) {
      // first slice in [s0, s1] with sbase[s] >= gfirst + tt * kSellQ1
      const uint64_t target = static_cast<uint64_t>(gfirst) + static_cast<uint64_t>(tt) * static_cast<uint64_t>(quantum);
      int64_t lo = s0, hi = s1;
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (static_cast<uint64_t>(sbase[mid]) < target) lo = mid + 1; else hi = mid;
      }
      return lo;
    };
    const int64_t tt = t - task_first_of_band[k];
    sell_task out;
    out.slice_first = static_cast<int32_t>(first_slice(tt));
    out.slice_end   = (t + 1 == task_first_of_band[k + 1]) ? static_cast<int32_t>(s1)
                                                           : static_cast<int32_t>(first_slice(tt + 1));
    out.g_first     = sbase[out.slice_first];
    out.n           = 0;
    if (out.slice_end > out.slice_first) {
      // group counts do not increase inside a band: first == last means all equal
      const uint32_t nf = sbase[out.slice_first + 1] - sbase[out.slice_first];
      const uint32_t nl = sbase[out.slice_end] - sbase[out.slice_end - 1];
      if (nf == nl) out.n = static_cast<int32_t>(nf);
    }
    tasks[t] = out;
  }
}

// ---- phase 2 --------------------------------------------------------------------------------
// One CTA per window of kSellWindow rows: stable counting sort of the window's rows by reduced
// length, descending.  Rows with more than kSellLong entries are appended to long_list (any
// order; the host sorts the few ids) and leave the window (row -1, length 0), like the rows
// behind the partition's last row.  Warp w owns the rows [128 w, 128 w + 128) of the window and
// ranks them in row order, so equal lengths keep their row order.
__global__ void __launch_bounds__(256) sell_window_sort_kernel(const int32_t* roff, int64_t nr,
                                                               int32_t* vrow, int32_t* vlen,
                                                               int32_t* long_list,
                                                               unsigned int* long_count,
                                                               unsigned int long_cap)
{
  constexpr int K = kSellLong + 1;   // keys 0 .. kSellLong
  static_assert(kSellWindow == 1024, "8 warps x 4 passes x 32 rows");
  __shared__ int hist[8][K + 1];
  __shared__ int start[K + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t w0 = int64_t(blockIdx.x) * kSellWindow;
  for (int i = tid; i < 8 * (K + 1); i += 256) (&hist[0][0])[i] = 0;
  __syncthreads();
  int key[4], len[4], row[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int64_t r = w0 + warp * 128 + p * 32 + lane;
    int l = 0, rr = -1;
    if (r < nr) {
      l  = roff[r + 1] - roff[r];
      rr = static_cast<int>(r);
      if (l > kSellLong) {
        const unsigned int j = atomicAdd(long_count, 1u);
        if (j < long_cap) long_list[j] = rr;
        l = 0; rr = -1;
      }
    }
    len[p] = l; row[p] = rr; key[p] = kSellLong - l;
    atomicAdd(&hist[warp][key[p]], 1);
  }
  __syncthreads();
  if (tid < K) {
    int run = 0;
    for (int w = 0; w < 8; ++w) { const int c = hist[w][tid]; hist[w][tid] = run; run += c; }
    start[tid] = run;   // bucket total for now
  }
  __syncthreads();
  if (warp == 0) {
    // exclusive scan of the K bucket totals by one warp (K = 129: five rounds of 32)
    int carry = 0;
    for (int b = 0; b < K; b += 32) {
      const int i = b + lane;
      int v = i < K ? start[i] : 0;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
      }
      if (i < K) start[i] = carry + inc - v;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
  }
  __syncthreads();
  const uint32_t lanes_lt = (1u << lane) - 1u;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const uint32_t m = __match_any_sync(0xffffffffu, key[p]);
    const int cur    = hist[warp][key[p]];
    const int pos    = start[key[p]] + cur + __popc(m & lanes_lt);
    __syncwarp();
    if ((m & lanes_lt) == 0) hist[warp][key[p]] = cur + __popc(m);
    __syncwarp();
    vrow[w0 + pos] = row[p];
    vlen[w0 + pos] = len[p];
  }
}

// skip_empty layouts: a row slice is kept when its first (longest) row has entries
__global__ void sell_keep_flags_kernel(const int32_t* vlen, int64_t nslices, uint32_t* keep)
{
  for (int64_t j = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; j < nslices;
       j += int64_t(gridDim.x) * blockDim.x)
    keep[j] = vlen[j * 32] > 0 ? 1u : 0u;
}
__global__ void sell_keep_map_kernel(const int32_t* vlen, const uint32_t* pos, int64_t nslices,
                                     int32_t* rmap)
{
  for (int64_t j = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; j < nslices;
       j += int64_t(gridDim.x) * blockDim.x)
    if (vlen[j * 32] > 0) rmap[pos[j]] = static_cast<int32_t>(j);
}

__global__ void sell_gather_lengths_kernel(const int32_t* roff, const int32_t* ids, int64_t n, int32_t* out)
{
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n;
       i += int64_t(gridDim.x) * blockDim.x)
    out[i] = roff[ids[i] + 1] - roff[ids[i]];
}

// length of every phase-2 slice: chunk slices = longest chunk in the slice, row slices = the
// length of the first (longest) row
__global__ void sell_p2_lengths_kernel(const int32_t* chunk_len, int64_t num_chunks,
                                       int64_t chunk_slices, const int32_t* vlen,
                                       const int32_t* rmap, int64_t nslices, uint32_t* slen)
{
  for (int64_t s = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; s < nslices;
       s += int64_t(gridDim.x) * blockDim.x) {
    if (s < chunk_slices) {
      int m = 0;
      for (int l = 0; l < 32; ++l) {
        const int64_t c = s * 32 + l;
        if (c < num_chunks && chunk_len[c] > m) m = chunk_len[c];
      }
      slen[s] = static_cast<uint32_t>(m);
    } else {
      const int64_t j = rmap ? rmap[s - chunk_slices] : s - chunk_slices;
      slen[s] = static_cast<uint32_t>(vlen[j * 32]);
    }
  }
}

// one warp per slice: lane l copies the entries of its row (or chunk) of the reduced CSC
template <typename RT>
__global__ void sell_p2_fill_kernel(const int32_t* roff, const int32_t* ridx, const RT* rw,
                                    const int32_t* vrow, const int32_t* rmap, const int32_t* chunk_row,
                                    const int32_t* chunk_off, const int32_t* chunk_len,
                                    int64_t num_chunks, int64_t chunk_slices, int64_t nslices,
                                    const uint32_t* sbase, int32_t dummy, int32_t x_split,
                                    const uint32_t* inv, int32_t* out_idx, RT* out_w, int32_t* out_row)
{
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarp = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t s = warp0; s < nslices; s += nwarp) {
    int64_t first = 0;
    int len = 0, row = -1;
    if (s < chunk_slices) {
      const int64_t c = s * 32 + lane;
      if (c < num_chunks) { first = int64_t(roff[chunk_row[c]]) + chunk_off[c]; len = chunk_len[c]; }
    } else {
      const int64_t j = rmap ? rmap[s - chunk_slices] : s - chunk_slices;
      row = vrow[j * 32 + lane];
      if (row >= 0) { first = roff[row]; len = roff[row + 1] - roff[row]; }
    }
    out_row[s * 32 + lane] = row;
    const uint32_t b0 = sbase[s], b1 = sbase[s + 1];
    for (uint32_t b = b0; b < b1; ++b) {
      const int j = static_cast<int>(b - b0);
      int32_t v = j < len ? ridx[first + j] : dummy;
      if (inv && j < len && v >= x_split) v = x_split + static_cast<int32_t>(inv[v - x_split]);
      out_idx[static_cast<int64_t>(b) * 32 + lane] = v;
      if (rw) out_w[static_cast<int64_t>(b) * 32 + lane] = j < len ? rw[first + j] : RT(0);
    }
  }
}

// task t covers the slices whose cost prefix (entries / 32 + 2 per slice) lies in
// [t * kSellQ2, (t + 1) * kSellQ2)
__global__ void sell_p2_tasks_kernel(const uint32_t* sbase, int64_t nslices, int64_t ntasks,
                                     int64_t quantum, int32_t* task_first)
{
  for (int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; t <= ntasks;
       t += int64_t(gridDim.x) * blockDim.x) {
    if (t == ntasks) { task_first[t] = static_cast<int32_t>(nslices); continue; }
    const uint64_t target = static_cast<uint64_t>(t) * static_cast<uint64_t>(quantum);
    int64_t lo = 0, hi = nslices;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (static_cast<uint64_t>(sbase[mid]) + 2ull * static_cast<uint64_t>(mid) < target) lo = mid + 1;
      else hi = mid;
    }
    task_first[t] = static_cast<int32_t>(lo);
  }
}

}  // namespace

namespace {
inline int env_int_or(const char* name, int dflt)
{
  const char* e = std::getenv(name);
  return e ? std::atoi(e) : dflt;
}
__global__ void sell_gather_u32_kernel(const uint32_t* src, const int64_t* at, int n, uint32_t* out)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    out[i] = src[at[i]];
}
}  // namespace

template <typename RT>
static void build_sell(b200_handle* h, edge_partition& ep, const RT* w, sell_layout& L, bool skip_empty)
{
  using G = sell_geom<RT>;
  cudaStream_t s   = h->stream;
  const int64_t ne = ep.num_edges;
  const int64_t nr = ep.num_rows;
  const int B      = G::entries;
  const int nb     = static_cast<int>((ep.num_minors + B - 1) / B);
  L.value_size     = static_cast<int>(sizeof(RT));
  L.band_entries   = B;
  L.enabled        = false;
  const char* env_min     = std::getenv("B200_BAND_MIN_EDGES");
  const int64_t min_edges = env_min ? std::atoll(env_min) : (int64_t{1} << 16);
  if (ne < std::max<int64_t>(min_edges, 64) || ne >= (int64_t{1} << 31) - 4096 || nb <= 0 ||
      nb > kSellMaxBands || ep.offsets64 || nr >= (int64_t{1} << 31) - 4096)
    return;
  CUDA_TRY(cudaStreamSynchronize(s));
  const auto t_begin = std::chrono::steady_clock::now();
  // cells are cut at CSC positions that are multiples of cell_max: long cells keep the partial
  // count of the hub rows down (large graphs), short ones keep a lane's longest walk - one task -
  // short enough for the CTAs of a small graph to finish together (profiles/r2_notes.md)
  const int cell_max = env_int_or("B200_SELL_CELL_MAX", ne >= (int64_t{1} << 29) ? kSellCellMax : 128);
  const int32_t* off = reinterpret_cast<const int32_t*>(ep.offsets.data());
  const int32_t* idx = ep.indices.data();
  const size_t n     = static_cast<size_t>(ne);

  // ---- 1-3: masks and scans -----------------------------------------------------------------
  dbuf<uint8_t> m8(n + 8, s);
  {
    dbuf<uint8_t> f8(n + 8, s), c8(n + 8, s), hot8(static_cast<size_t>(nb), s);
    dbuf<unsigned long long> hist(static_cast<size_t>(nb), s);
    CUDA_TRY(cudaMemsetAsync(f8.data(), 0, n + 8, s));
    CUDA_TRY(cudaMemsetAsync(hist.data(), 0, sizeof(unsigned long long) * nb, s));
    sell_row_starts_kernel<<<sgrid(static_cast<size_t>(nr)), kBT, 0, s>>>(off, nr, f8.data());
    sell_cell_starts_kernel<<<sgrid(n), kBT, 0, s>>>(idx, ne, B, cell_max, f8.data());
    sell_classify_kernel<<<sgrid(n), kBT, static_cast<size_t>(nb) * 4, s>>>(f8.data(), idx, ne, B, nb,
                                                                            c8.data(), hist.data());
    const char* env = std::getenv("B200_BAND_HOT_MIN");
    const long long hot_min = env ? std::atoll(env) : kHotBandMin;
    sell_mark_hot_kernel<<<sgrid(static_cast<size_t>(nb)), kBT, 0, s>>>(hist.data(), nb, hot_min,
                                                                        hot8.data());
    sell_masks_kernel<<<sgrid(n), kBT, 0, s>>>(f8.data(), c8.data(), hot8.data(), idx, ne, B, m8.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  dbuf<uint32_t> newpos(n + 1, s), pid(n + 1, s);
  exclusive_scan<s8_keep, uint32_t>(reinterpret_cast<const s8_keep*>(m8.data()), newpos.data(), n, s, true);
  exclusive_scan<s8_head, uint32_t>(reinterpret_cast<const s8_head*>(m8.data()), pid.data(), n, s, true);
  uint32_t totals[2] = {0, 0};
  CUDA_TRY(cudaMemcpyAsync(&totals[0], newpos.data() + n, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&totals[1], pid.data() + n, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  const int64_t n_red = totals[0], n_part = totals[1];
  const int64_t n_band = ne - n_red + n_part;
  if (n_band < ne / 8 || n_part == 0 || ep.num_minors + n_part + 2 >= (int64_t{1} << 31) - 1) return;
  L.num_partials     = n_part;
  L.num_banded_edges = n_band;
  L.x_split          = static_cast<int32_t>(ep.num_minors);
  L.direct_entries   = n_red - n_part;

  // ---- reduced CSC (temporary) --------------------------------------------------------------
  dbuf<int32_t> roff(static_cast<size_t>(nr) + 1, s), ridx(static_cast<size_t>(n_red), s, 16);
  dbuf<RT> rw;
  if (w) rw.alloc(static_cast<size_t>(n_red), s, 16);
  sell_fill_reduced_kernel<RT><<<sgrid(n), kBT, 0, s>>>(m8.data(), newpos.data(), pid.data(), idx, w, ne,
                                                        L.x_split, ridx.data(), w ? rw.data() : nullptr);
  sell_reduced_offsets_kernel<<<sgrid(static_cast<size_t>(nr) + 1), kBT, 0, s>>>(off, nr + 1, newpos.data(),
                                                                                 roff.data());
  CUDA_TRY(cudaGetLastError());

  // ---- phase 1: cells sorted by (band, length descending) -----------------------------------
  const size_t np = static_cast<size_t>(n_part);
  // Partial slots: row-major (a row's partials are consecutive: phase 2 reads them with few
  // sectors, phase 1 scatters its stores) while the partial vector stays L2-resident; beyond
  // that every scattered store is a read-modify-write of a DRAM sector, and the slots follow
  // the sorted cell order instead (phase 1 stores coalesced, phase 2 gathers).
  const char* env_sb       = std::getenv("B200_SELL_SLOT_BYTES");
  const int64_t slot_bytes = env_sb ? std::atoll(env_sb) : 0;   // measured: band-major wins at every size (profiles/r2_notes.md)
  const bool band_major    = static_cast<int64_t>(np) * static_cast<int64_t>(sizeof(RT)) > slot_bytes;
  L.band_major_slots       = band_major;
  dbuf<uint32_t> inv;
  if (band_major) inv.alloc(np, s);
  dbuf<uint32_t> ce0(np, s), v0(np, s), v1(np, s);
  dbuf<uint64_t> k0(np, s), k1(np, s);
  dbuf<unsigned long long> total(1, s);
  CUDA_TRY(cudaMemsetAsync(total.data(), 0, 8, s));
  sell_cell_records_kernel<<<sgrid(n), kBT, 0, s>>>(m8.data(), pid.data(), idx, ne, B, cell_max, ce0.data(),
                                                    k0.data(), v0.data(), total.data());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  newpos.release();
  pid.release();
  m8.release();
  const int band_key_bits = bits_needed(static_cast<uint64_t>(nb - 1));
  const int which = radix_sort<uint64_t>(k0.data(), k1.data(), v0.data(), v1.data(), np,
                                         {{0, 10}, {32, 32 + band_key_bits}}, s);
  const uint64_t* keys = which ? k1.data() : k0.data();
  const uint32_t* vals = which ? v1.data() : v0.data();
  dbuf<int64_t> d_bstart(static_cast<size_t>(nb) + 1, s), d_slice_first(static_cast<size_t>(nb) + 1, s);
  sell_band_starts_kernel<<<sgrid(static_cast<size_t>(nb) + 1), kBT, 0, s>>>(keys, n_part, nb, d_bstart.data());
  CUDA_TRY(cudaGetLastError());
  std::vector<int64_t> bstart(static_cast<size_t>(nb) + 1), slice_first(static_cast<size_t>(nb) + 1);
  unsigned long long banded_edges = 0;
  CUDA_TRY(cudaMemcpyAsync(bstart.data(), d_bstart.data(), bstart.size() * 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&banded_edges, total.data(), 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(static_cast<int64_t>(banded_edges) == n_band, kUnknownError,
               "sell layout: cell lengths do not add up to the pre-reduced edges");
  slice_first[0] = 0;
  for (int k = 0; k < nb; ++k) slice_first[k + 1] = slice_first[k] + div_up(bstart[k + 1] - bstart[k], 32);
  const int64_t S1 = slice_first[nb];
  B200_EXPECTS(S1 > 0 && S1 < (int64_t{1} << 31) - 1, kUnknownError, "sell layout: no phase-1 slices");
  CUDA_TRY(cudaMemcpyAsync(d_slice_first.data(), slice_first.data(), slice_first.size() * 8,
                           cudaMemcpyHostToDevice, s));
  L.p1_slices = S1;
  L.p1_sbase.alloc(static_cast<size_t>(S1) + 1, s);
  sell_p1_groups_kernel<<<sgrid(static_cast<size_t>(S1)), kBT, 0, s>>>(keys, d_bstart.data(),
                                                                      d_slice_first.data(), nb, S1, cell_max,
                                                                      L.p1_sbase.data());
  CUDA_TRY(cudaGetLastError());
  exclusive_scan<uint32_t, uint32_t>(L.p1_sbase.data(), L.p1_sbase.data(), static_cast<size_t>(S1), s, true);
  uint32_t p1_groups = 0;
  CUDA_TRY(cudaMemcpyAsync(&p1_groups, L.p1_sbase.data() + S1, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  L.p1_groups = p1_groups;
  const size_t p1_stored = static_cast<size_t>(p1_groups) * 32 * kSellG;
  L.p1_idx.alloc(p1_stored, s, 64);
  if (w) L.p1_w.alloc(p1_stored * sizeof(RT), s, 64);
  L.p1_slot.alloc(static_cast<size_t>(S1) * 32, s);
  sell_p1_fill_kernel<RT><<<sgrid(static_cast<size_t>(S1) * 32), kBT, 0, s>>>(
    keys, vals, ce0.data(), d_bstart.data(), d_slice_first.data(), nb, S1, L.p1_sbase.data(), idx, w, B,
    cell_max, L.p1_idx.data(), w ? reinterpret_cast<RT*>(L.p1_w.data()) : nullptr, L.p1_slot.data(),
    band_major ? inv.data() : nullptr);
  CUDA_TRY(cudaGetLastError());
  // tasks and units
  {
    dbuf<uint32_t> d_g(static_cast<size_t>(nb) + 1, s);
    sell_gather_u32_kernel<<<sgrid(static_cast<size_t>(nb) + 1), kBT, 0, s>>>(
      L.p1_sbase.data(), d_slice_first.data(), nb + 1, d_g.data());
    CUDA_TRY(cudaGetLastError());
    std::vector<uint32_t> gfirst(static_cast<size_t>(nb) + 1);
    CUDA_TRY(cudaMemcpyAsync(gfirst.data(), d_g.data(), gfirst.size() * 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    std::vector<int64_t> task_first(static_cast<size_t>(nb) + 1);
    task_first[0] = 0;
    // small graphs: at least ~4 tasks for every warp of the grid
    const int64_t q1 = std::max<int64_t>(4, std::min<int64_t>(kSellQ1, static_cast<int64_t>(p1_groups) /
                                                                 (int64_t{h->sm_count} * 32 * 4)));
    for (int k = 0; k < nb; ++k)
      task_first[k + 1] = task_first[k] + div_up(static_cast<int64_t>(gfirst[k + 1] - gfirst[k]), q1);
    const int64_t T1 = task_first[nb];
    B200_EXPECTS(T1 > 0 && T1 < (int64_t{1} << 31), kUnknownError, "sell layout: no phase-1 tasks");
    dbuf<int64_t> d_task_first(static_cast<size_t>(nb) + 1, s);
    CUDA_TRY(cudaMemcpyAsync(d_task_first.data(), task_first.data(), task_first.size() * 8,
                             cudaMemcpyHostToDevice, s));
    L.p1_tasks.alloc(static_cast<size_t>(T1), s);
    L.p1_num_tasks = T1;
    sell_p1_tasks_kernel<<<sgrid(static_cast<size_t>(T1)), kBT, 0, s>>>(
      d_task_first.data(), d_slice_first.data(), nb, L.p1_sbase.data(), T1, q1, L.p1_tasks.data());
    CUDA_TRY(cudaGetLastError());
    // units of consecutive tasks of one band, in band order; sizes follow guided self-scheduling
    const int nctas = h->sm_count;
    const int64_t kUnitMin = 32, kUnitMax = 1024;
    std::vector<band_group> units;
    int band = 0;
    for (int64_t k = 0; k < T1;) {
      while (task_first[band + 1] <= k) ++band;
      const int64_t size = std::max<int64_t>(kUnitMin, std::min<int64_t>(kUnitMax, (T1 - k) / (2 * nctas)));
      const int64_t e    = std::min<int64_t>({T1, k + size, task_first[band + 1]});
      band_group u{};
      u.task_first = static_cast<int32_t>(k);
      u.task_end   = static_cast<int32_t>(e);
      u.band       = band;
      units.push_back(u);
      k = e;
    }
    L.p1_num_units = static_cast<int64_t>(units.size());
    L.p1_unit_band.resize(units.size());
    for (size_t k = 0; k < units.size(); ++k) L.p1_unit_band[k] = units[k].band;
    L.p1_units.alloc(units.size(), s);
    CUDA_TRY(cudaMemcpyAsync(L.p1_units.data(), units.data(), units.size() * sizeof(band_group),
                             cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  ce0.release(); v0.release(); v1.release(); k0.release(); k1.release();

  // ---- phase 2: rows of the reduced CSC, window-sorted by length ----------------------------
  const int64_t windows = div_up(nr, kSellWindow);
  dbuf<int32_t> vrow(static_cast<size_t>(windows) * kSellWindow, s), vlen(static_cast<size_t>(windows) * kSellWindow, s);
  const unsigned int long_cap = static_cast<unsigned int>(n_red / kSellLong + 8);
  dbuf<int32_t> long_list(long_cap, s);
  dbuf<unsigned int> long_count(1, s);
  CUDA_TRY(cudaMemsetAsync(long_count.data(), 0, 4, s));
  sell_window_sort_kernel<<<static_cast<unsigned>(windows), 256, 0, s>>>(
    roff.data(), nr, vrow.data(), vlen.data(), long_list.data(), long_count.data(), long_cap);
  CUDA_TRY(cudaGetLastError());
  unsigned int n_long = 0;
  CUDA_TRY(cudaMemcpyAsync(&n_long, long_count.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(n_long <= long_cap, kUnknownError, "sell layout: long-row list overflow");
  std::vector<int32_t> long_ids(n_long), long_len(n_long);
  std::vector<int32_t> chunk_row, chunk_off, chunk_len;
  std::vector<int64_t> chunk_first(static_cast<size_t>(n_long) + 1, 0);
  if (n_long) {
    CUDA_TRY(cudaMemcpyAsync(long_ids.data(), long_list.data(), size_t(n_long) * 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    std::sort(long_ids.begin(), long_ids.end());
    CUDA_TRY(cudaMemcpyAsync(long_list.data(), long_ids.data(), size_t(n_long) * 4, cudaMemcpyHostToDevice, s));
    dbuf<int32_t> d_len(n_long, s);
    sell_gather_lengths_kernel<<<sgrid(n_long), kBT, 0, s>>>(roff.data(), long_list.data(), n_long, d_len.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(long_len.data(), d_len.data(), size_t(n_long) * 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    for (unsigned int i = 0; i < n_long; ++i) {
      const int32_t len = long_len[i];
      for (int32_t o = 0; o < len; o += kSellLong) {
        chunk_row.push_back(long_ids[i]);
        chunk_off.push_back(o);
        chunk_len.push_back(std::min<int32_t>(kSellLong, len - o));
      }
      chunk_first[i + 1] = static_cast<int64_t>(chunk_row.size());
    }
  }
  const int64_t num_chunks = static_cast<int64_t>(chunk_row.size());
  // skip_empty (multi-GPU partial sums: the caller clears the output): slices whose rows all
  // have no entry are left out - most rows of a 2D partition have no edge in its column block
  int64_t RS = windows * (kSellWindow / 32);
  dbuf<int32_t> rmap;
  L.skips_empty_rows = false;
  if (skip_empty) {
    dbuf<uint32_t> keep(static_cast<size_t>(RS) + 1, s);
    sell_keep_flags_kernel<<<sgrid(static_cast<size_t>(RS)), kBT, 0, s>>>(vlen.data(), RS, keep.data());
    CUDA_TRY(cudaGetLastError());
    exclusive_scan<uint32_t, uint32_t>(keep.data(), keep.data(), static_cast<size_t>(RS), s, true);
    uint32_t kept = 0;
    CUDA_TRY(cudaMemcpyAsync(&kept, keep.data() + RS, 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    rmap.alloc(std::max<size_t>(kept, 1), s);
    sell_keep_map_kernel<<<sgrid(static_cast<size_t>(RS)), kBT, 0, s>>>(vlen.data(), keep.data(), RS, rmap.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));
    RS = kept;
    L.skips_empty_rows = true;
  }
  const int32_t* rmap_p = skip_empty ? rmap.data() : nullptr;
  const int64_t CS = div_up(num_chunks, 32), S2 = CS + RS;
  B200_EXPECTS(S2 < (int64_t{1} << 31) - 1, kUnknownError, "sell layout: too many phase-2 slices");
  L.num_long_rows   = n_long;
  L.num_chunks      = num_chunks;
  L.p2_chunk_slices = CS;
  L.p2_slices       = S2;
  L.long_row_ids.alloc(std::max<size_t>(n_long, 1), s);
  L.long_row_chunk_first.alloc(static_cast<size_t>(n_long) + 1, s);
  dbuf<int32_t> d_crow(std::max<size_t>(chunk_row.size(), 1), s), d_coff(std::max<size_t>(chunk_row.size(), 1), s),
    d_clen(std::max<size_t>(chunk_row.size(), 1), s);
  if (n_long) {
    CUDA_TRY(cudaMemcpyAsync(L.long_row_ids.data(), long_ids.data(), size_t(n_long) * 4, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_crow.data(), chunk_row.data(), chunk_row.size() * 4, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_coff.data(), chunk_off.data(), chunk_off.size() * 4, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_clen.data(), chunk_len.data(), chunk_len.size() * 4, cudaMemcpyHostToDevice, s));
  }
  CUDA_TRY(cudaMemcpyAsync(L.long_row_chunk_first.data(), chunk_first.data(), chunk_first.size() * 8,
                           cudaMemcpyHostToDevice, s));
  L.p2_sbase.alloc(static_cast<size_t>(S2) + 1, s);
  sell_p2_lengths_kernel<<<sgrid(static_cast<size_t>(S2)), kBT, 0, s>>>(d_clen.data(), num_chunks, CS,
                                                                       vlen.data(), rmap_p, S2, L.p2_sbase.data());
  CUDA_TRY(cudaGetLastError());
  exclusive_scan<uint32_t, uint32_t>(L.p2_sbase.data(), L.p2_sbase.data(), static_cast<size_t>(S2), s, true);
  uint32_t p2_rows = 0;
  CUDA_TRY(cudaMemcpyAsync(&p2_rows, L.p2_sbase.data() + S2, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  L.p2_entries = static_cast<int64_t>(p2_rows) * 32;
  L.p2_idx.alloc(static_cast<size_t>(L.p2_entries), s, 64);
  if (w) L.p2_w.alloc(static_cast<size_t>(L.p2_entries) * sizeof(RT), s, 64);
  L.p2_row.alloc(static_cast<size_t>(S2) * 32, s);
  const int32_t dummy = static_cast<int32_t>(L.x_split + n_part);   // pval[n_part] stays zero
  sell_p2_fill_kernel<RT><<<sgrid(static_cast<size_t>(S2) * 32), kBT, 0, s>>>(
    roff.data(), ridx.data(), w ? rw.data() : nullptr, vrow.data(), rmap_p, d_crow.data(), d_coff.data(),
    d_clen.data(), num_chunks, CS, S2, L.p2_sbase.data(), dummy, L.x_split,
    band_major ? inv.data() : nullptr, L.p2_idx.data(),
    w ? reinterpret_cast<RT*>(L.p2_w.data()) : nullptr, L.p2_row.data());
  CUDA_TRY(cudaGetLastError());
  const int64_t cost = static_cast<int64_t>(p2_rows) + 2 * S2;
  // small graphs: at least ~4 tasks for every warp the kernel can have in flight
  const int64_t q2   = std::max<int64_t>(8, std::min<int64_t>(kSellQ2, cost / (int64_t{h->sm_count} * 32 * 4)));
  const int64_t T2   = std::max<int64_t>(1, div_up(cost, q2));
  B200_EXPECTS(T2 < (int64_t{1} << 31) - 1, kUnknownError, "sell layout: too many phase-2 tasks");
  L.p2_num_tasks = T2;
  L.p2_task_first.alloc(static_cast<size_t>(T2) + 1, s);
  sell_p2_tasks_kernel<<<sgrid(static_cast<size_t>(T2) + 1), kBT, 0, s>>>(L.p2_sbase.data(), S2, T2, q2,
                                                                         L.p2_task_first.data());
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  L.enabled  = true;
  L.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
  if (std::getenv("B200_BAND_VERBOSE")) {
    fprintf(stderr,
            "[b200] sell layout: E=%lld band=%d entries, pre-reduced=%lld (%.3f) cells=%lld (%.4f) "
            "p1 stored=%lld (%.3f) slices=%lld tasks=%lld units=%lld | direct=%lld (%.3f) p2 stored=%lld "
            "(%.3f) slices=%lld long rows=%lld chunks=%lld tasks=%lld slots=%s build=%.1f ms\n",
            (long long)ne, B, (long long)n_band, double(n_band) / double(ne), (long long)n_part,
            double(n_part) / double(ne), (long long)p1_stored, double(p1_stored) / double(ne),
            (long long)S1, (long long)L.p1_num_tasks, (long long)L.p1_num_units,
            (long long)L.direct_entries, double(L.direct_entries) / double(ne), (long long)L.p2_entries,
            double(L.p2_entries) / double(ne), (long long)S2, (long long)n_long, (long long)num_chunks,
            (long long)T2, band_major ? "band-major" : "row-major", L.build_ms);
  }
}

template <typename RT>
sell_layout* sell_get(b200_handle* h, edge_partition& ep, const RT* weights, bool skip_empty)
{
  const int slot = sizeof(RT) == 4 ? 0 : 1;
  if (!ep.sell[slot]) {
    auto L = std::make_shared<sell_layout>();
    build_sell<RT>(h, ep, weights, *L, skip_empty);
    ep.sell[slot] = L;
  }
  return ep.sell[slot].get();
}

template sell_layout* sell_get<float>(b200_handle*, edge_partition&, const float*, bool);
template sell_layout* sell_get<double>(b200_handle*, edge_partition&, const double*, bool);

}  // namespace b200
