// Build of the banded pre-reduction layout (banded.cuh) from a CSC rectangle, on the GPU with
// the hand-written scan / radix sort of prims.cuh.  Single GPU, 32-bit offsets.
//
//   1. cell starts:  edge e starts a cell if it is the first edge of its row, or its source
//      lies in another band than the source of edge e - 1, or e is a multiple of kCellMax
//      (bounds the cell size; any cut is valid because phase 2 sums all partials of a row);
//   2. a cell is a CANDIDATE if it has >= kMinCell edges (decided per edge from the cell-start
//      flags of the 2 * kMinCell - 1 edges around it); a band is HOT if its candidate cells
//      hold >= kHotBandMin edges; the cells of hot bands are pre-reduced ("banded");
//   3. three scans give every edge its position in the reduced CSC (kept edges: all direct
//      edges + the first edge of every banded cell), the slot of its cell's partial sum
//      (row-major, so that a row's partials are consecutive) and its rank among the banded
//      edges; the reduced CSC is written in one pass;
//   4. the banded edges are sorted by band with the stable LSD radix sort (1-2 passes over the
//      band bits of (band << 32 | edge) keys), which keeps every cell contiguous;
//   5. band-major arrays: 16-bit band-local index | cell-start flag, weights, and for every
//      cell (in band-major order) the slot of its partial;
//   6. tasks = the cells that start inside one window of kTaskWindow edges of a band; the host
//      cuts the task list into the units of the kernel's work queue.
#include <chrono>
#include <cstdlib>

#include "banded.cuh"
#include "graph.cuh"

namespace b200 {

namespace {

constexpr int kBT = 256;
inline unsigned bgrid(size_t n, size_t cap = 148 * 32)
{
  size_t b = (n + kBT - 1) / kBT;
  if (b > cap) b = cap;
  if (b == 0) b = 1;
  return static_cast<unsigned>(b);
}

// per-edge mask byte: bit 0 keep (entry of the reduced CSC), bit 1 head (first edge of a banded
// cell), bit 2 banded, bit 3 cell start.  The scans read single bits through these wrappers.
struct m8_keep { uint8_t v; __host__ __device__ operator uint32_t() const { return v & 1u; } };
struct m8_head { uint8_t v; __host__ __device__ operator uint32_t() const { return (v >> 1) & 1u; } };
struct m8_band { uint8_t v; __host__ __device__ operator uint32_t() const { return (v >> 2) & 1u; } };
struct u8_one  { uint8_t v; __host__ __device__ operator uint32_t() const { return v; } };

__global__ void mark_row_starts_kernel(const int32_t* off, int64_t n_rows, uint8_t* f8)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < n_rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int32_t b = off[r], e = off[r + 1];
    if (b < e) f8[b] = 1;
  }
}

// f8[e]: bit 0 row start (input), bit 1 cell start (output)
__global__ void mark_cell_starts_kernel(const int32_t* idx, int64_t ne, int bits, uint8_t* f8)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const uint8_t rs = f8[e] & 1u;
    bool f = rs || e == 0 || (e % kCellMax) == 0;
    if (!f) f = (idx[e] >> bits) != (idx[e - 1] >> bits);
    f8[e] = static_cast<uint8_t>(rs | (f ? 2u : 0u));
  }
}

// c8[e] = 1 if the cell of edge e has >= kMinCell edges; hist[band] += 1 for those edges.
// The histogram is kept per block in shared memory (nb <= kMaxBands bins); inside a warp the
// 32 consecutive edges form a few runs of equal band (neighbour lists ascend), and only the
// first lane of a run adds the run's candidate count.
constexpr int kMaxBands = 8192;

__global__ void classify_kernel(const uint8_t* f8, const int32_t* idx, int64_t ne, int bits, int nb,
                                uint8_t* c8, unsigned long long* hist)
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
      band = idx[e] >> bits;
      // nearest cell start at or before e, within kMinCell - 1 edges
      int64_t p = -1;
#pragma unroll
      for (int j = 0; j < kMinCell; ++j) {
        if (p < 0 && e - j >= 0 && (f8[e - j] & 2u)) p = e - j;
      }
      cand = true;  // no start found: the cell started >= kMinCell edges ago
      if (p >= 0) {
#pragma unroll
        for (int j = 1; j < kMinCell; ++j) {
          const int64_t q = p + j;
          if (q >= ne || (f8[q] & 2u)) cand = false;
        }
      }
      c8[e] = cand ? 1 : 0;
    }
    // runs of equal band inside the warp
    const int prev       = __shfl_up_sync(0xffffffffu, band, 1);
    const bool head      = lane == 0 || prev != band;
    const uint32_t heads = __ballot_sync(0xffffffffu, head);
    const uint32_t cands = __ballot_sync(0xffffffffu, cand);
    if (head && band >= 0) {
      const uint32_t above = heads & ~((2u << lane) - 1u);          // heads after this lane
      const int end        = above ? __ffs(above) - 1 : 32;         // first lane of the next run
      const uint32_t run   = (end >= 32 ? 0xffffffffu : ((1u << end) - 1u)) & ~((1u << lane) - 1u);
      const int n          = __popc(cands & run);
      if (n) atomicAdd(&s_hist[band], static_cast<uint32_t>(n));
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nb; k += blockDim.x)
    if (s_hist[k]) atomicAdd(hist + k, static_cast<unsigned long long>(s_hist[k]));
}

__global__ void mark_hot_kernel(const unsigned long long* hist, int nb, long long min_edges,
                                uint8_t* hot8)
{
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < nb; k += gridDim.x * blockDim.x)
    hot8[k] = hist[k] >= static_cast<unsigned long long>(min_edges) ? 1 : 0;
}

__global__ void finalize_masks_kernel(const uint8_t* f8, const uint8_t* c8, const uint8_t* hot8,
                                      const int32_t* idx, int64_t ne, int bits, uint8_t* m8)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const bool f      = (f8[e] & 2u) != 0;
    const bool banded = c8[e] && hot8[idx[e] >> bits];
    const bool keep   = !banded || f;
    const bool head   = banded && f;
    m8[e] = static_cast<uint8_t>((keep ? 1u : 0u) | (head ? 2u : 0u) | (banded ? 4u : 0u) |
                                 (f ? 8u : 0u));
  }
}

template <typename RT>
__global__ void fill_reduced_kernel(const uint8_t* m8, const uint32_t* newpos, const uint32_t* pid,
                                    const uint32_t* bpos, const int32_t* idx, const RT* w,
                                    int64_t ne, int bits, int32_t x_split, int32_t* ridx, RT* rw,
                                    uint64_t* keys)
{
  for (int64_t e = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; e < ne;
       e += int64_t(gridDim.x) * blockDim.x) {
    const uint8_t m   = m8[e];
    const bool banded = (m & 4u) != 0;
    if (m & 1u) {
      const uint32_t j = newpos[e];
      ridx[j] = banded ? static_cast<int32_t>(x_split + static_cast<int32_t>(pid[e])) : idx[e];
      if (w) rw[j] = banded ? RT(1) : w[e];
    }
    if (banded)
      keys[bpos[e]] = (static_cast<uint64_t>(static_cast<uint32_t>(idx[e] >> bits)) << 32) |
                      static_cast<uint64_t>(e);
  }
}

__global__ void reduced_offsets_kernel(const int32_t* off, int64_t n, const uint32_t* newpos,
                                       int32_t* roff)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < n;
       r += int64_t(gridDim.x) * blockDim.x)
    roff[r] = static_cast<int32_t>(newpos[off[r]]);
}

template <typename RT>
__global__ void fill_band_major_kernel(const uint64_t* keys, int64_t nb_edges, const int32_t* idx,
                                       const uint8_t* m8, const RT* w, int bits, uint16_t* bidx,
                                       RT* bw, uint8_t* hf8)
{
  const uint32_t mask = (1u << bits) - 1u;
  for (int64_t j = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; j < nb_edges;
       j += int64_t(gridDim.x) * blockDim.x) {
    const uint32_t e = static_cast<uint32_t>(keys[j] & 0xffffffffu);
    const bool f     = (m8[e] & 8u) != 0;
    bidx[j] = static_cast<uint16_t>((static_cast<uint32_t>(idx[e]) & mask) | (f ? 0x8000u : 0u));
    if (w) bw[j] = w[e];
    hf8[j] = f ? 1 : 0;
  }
}

__global__ void fill_ppos_kernel(const uint64_t* keys, const uint8_t* hf8, const uint32_t* rank,
                                 const uint32_t* pid, int64_t nb_edges, int32_t* ppos,
                                 uint32_t* fpos)
{
  for (int64_t j = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; j < nb_edges;
       j += int64_t(gridDim.x) * blockDim.x) {
    if (hf8[j]) {
      const uint32_t e = static_cast<uint32_t>(keys[j] & 0xffffffffu);
      const uint32_t q = rank[j];
      ppos[q] = static_cast<int32_t>(pid[e]);
      fpos[q] = static_cast<uint32_t>(j);
    }
  }
}

// bstart[k] = first band-major position whose band is >= k, k in [0, nb]
__global__ void band_starts_kernel(const uint64_t* keys, int64_t nb_edges, int nb, int64_t* bstart)
{
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= nb; k += gridDim.x * blockDim.x) {
    const uint64_t target = static_cast<uint64_t>(k) << 32;
    int64_t lo = 0, hi = nb_edges;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < target) lo = mid + 1; else hi = mid;
    }
    bstart[k] = lo;
  }
}

// task k covers the cells that start in the window [win_lo[k], win_hi[k]) of band-major
// positions (win_hi is the next window's start or the band's end, both of which are followed
// by a cell start or the end of the array)
__global__ void finalize_tasks_kernel(const int64_t* win_lo, const int64_t* win_hi,
                                      const int32_t* win_band, int64_t ntasks, const uint32_t* rank,
                                      const uint32_t* fpos, int64_t nb_edges, int64_t npartials,
                                      band_task* tasks)
{
  for (int64_t k = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; k < ntasks;
       k += int64_t(gridDim.x) * blockDim.x) {
    auto next_start = [&](int64_t p, int64_t* q_out) {
      // first cell start at or after band-major position p
      if (p >= nb_edges) { *q_out = npartials; return nb_edges; }
      const int64_t q = rank[p];
      *q_out          = q;
      return q < npartials ? static_cast<int64_t>(fpos[q]) : nb_edges;
    };
    int64_t q0, q1;
    const int64_t a = next_start(win_lo[k], &q0);
    const int64_t b = next_start(win_hi[k], &q1);
    band_task t;
    t.first      = a;
    t.rank_first = q0;
    t.count      = static_cast<int32_t>(b > a ? b - a : 0);
    t.band       = win_band[k];
    tasks[k]     = t;
  }
}

}  // namespace

template <typename RT>
static void build_banded(b200_handle* h, edge_partition& ep, const RT* w, banded_layout& L,
                         int tile_edges, int row_cap)
{
  using G = band_geom<RT>;
  cudaStream_t s   = h->stream;
  const int64_t ne = ep.num_edges;
  const int64_t nr = ep.num_rows;
  const int bits   = G::bits;
  const int nb     = static_cast<int>((ep.num_minors + G::entries - 1) >> bits);
  L.band_bits      = bits;
  L.value_size     = static_cast<int>(sizeof(RT));
  L.enabled        = false;
  const char* env_min   = std::getenv("B200_BAND_MIN_EDGES");
  const int64_t min_edges = env_min ? std::atoll(env_min) : (int64_t{1} << 16);
  if (ne < std::max<int64_t>(min_edges, 64) || ne >= (int64_t{1} << 31) - 4096 || nb <= 0 ||
      nb > kMaxBands || ep.offsets64 || !ep.has_segments)
    return;
  CUDA_TRY(cudaStreamSynchronize(s));
  const auto t_begin = std::chrono::steady_clock::now();
  const int32_t* off = reinterpret_cast<const int32_t*>(ep.offsets.data());
  const int32_t* idx = ep.indices.data();
  const size_t n     = static_cast<size_t>(ne);

  dbuf<uint8_t> f8(n + 8, s), m8(n + 8, s);
  dbuf<uint8_t> hot8(static_cast<size_t>(nb), s);
  dbuf<unsigned long long> hist(static_cast<size_t>(nb), s);
  CUDA_TRY(cudaMemsetAsync(f8.data(), 0, n + 8, s));
  CUDA_TRY(cudaMemsetAsync(hist.data(), 0, sizeof(unsigned long long) * nb, s));
  mark_row_starts_kernel<<<bgrid(static_cast<size_t>(nr)), kBT, 0, s>>>(off, nr, f8.data());
  mark_cell_starts_kernel<<<bgrid(n), kBT, 0, s>>>(idx, ne, bits, f8.data());
  {
    dbuf<uint8_t> c8(n + 8, s);
    classify_kernel<<<bgrid(n), kBT, static_cast<size_t>(nb) * 4, s>>>(f8.data(), idx, ne, bits, nb,
                                                                       c8.data(), hist.data());
    const char* env = std::getenv("B200_BAND_HOT_MIN");
    const long long hot_min = env ? std::atoll(env) : kHotBandMin;
    mark_hot_kernel<<<bgrid(static_cast<size_t>(nb)), kBT, 0, s>>>(hist.data(), nb, hot_min,
                                                                   hot8.data());
    finalize_masks_kernel<<<bgrid(n), kBT, 0, s>>>(f8.data(), c8.data(), hot8.data(), idx, ne, bits,
                                                   m8.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  f8.release();

  dbuf<uint32_t> newpos(n + 1, s), pid(n + 1, s), bpos(n + 1, s);
  exclusive_scan<m8_keep, uint32_t>(reinterpret_cast<const m8_keep*>(m8.data()), newpos.data(), n, s,
                                    true);
  exclusive_scan<m8_head, uint32_t>(reinterpret_cast<const m8_head*>(m8.data()), pid.data(), n, s,
                                    true);
  exclusive_scan<m8_band, uint32_t>(reinterpret_cast<const m8_band*>(m8.data()), bpos.data(), n, s,
                                    true);
  uint32_t totals[3] = {0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(&totals[0], newpos.data() + n, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&totals[1], pid.data() + n, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&totals[2], bpos.data() + n, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  const int64_t n_red = totals[0], n_part = totals[1], n_band = totals[2];
  // not worth a second kernel when little can be pre-reduced
  if (n_band < ne / 8 || n_part == 0 || ep.num_minors + n_part >= (int64_t{1} << 31) - 1) return;

  L.num_banded_edges = n_band;
  L.num_partials     = n_part;
  L.x_split          = static_cast<int32_t>(ep.num_minors);

  // ---- reduced CSC ------------------------------------------------------------------------
  L.reduced = std::make_unique<edge_partition>();
  edge_partition& rp = *L.reduced;
  rp.num_rows    = nr;
  rp.num_edges   = n_red;
  rp.major_first = ep.major_first;
  rp.minor_first = ep.minor_first;
  rp.num_minors  = ep.num_minors;
  rp.offsets64   = false;
  rp.offsets.alloc((static_cast<size_t>(nr) + 1) * 4, s, 64);
  rp.indices.alloc(static_cast<size_t>(n_red), s, 16);
  if (w) rp.weights.alloc(static_cast<size_t>(n_red) * sizeof(RT), s, 32);
  for (int i = 0; i < 5; ++i) rp.seg[i] = ep.seg[i];
  rp.has_segments        = true;
  rp.generic_schedule    = true;   // long rows are listed by id (tail kernel)
  L.generic_rows         = ep.generic_schedule;
  rp.minor_piece_offsets = {0, ep.num_minors};

  dbuf<uint64_t> k0(static_cast<size_t>(n_band), s), k1(static_cast<size_t>(n_band), s);
  fill_reduced_kernel<RT><<<bgrid(n), kBT, 0, s>>>(
    m8.data(), newpos.data(), pid.data(), bpos.data(), idx, w, ne, bits, L.x_split,
    rp.indices.data(), w ? reinterpret_cast<RT*>(rp.weights.data()) : nullptr, k0.data());
  reduced_offsets_kernel<<<bgrid(static_cast<size_t>(nr) + 1), kBT, 0, s>>>(
    off, nr + 1, newpos.data(), reinterpret_cast<int32_t*>(rp.offsets.data()));
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(s));
  newpos.release();
  bpos.release();

  // ---- band-major arrays ------------------------------------------------------------------
  const int band_key_bits = bits_needed(static_cast<uint64_t>(nb - 1));
  const int which = radix_sort<uint64_t>(k0.data(), k1.data(), nullptr, nullptr,
                                         static_cast<size_t>(n_band),
                                         {{32, 32 + band_key_bits}}, s);
  const uint64_t* keys = which ? k1.data() : k0.data();
  constexpr size_t kPad = 1024;  // a warp may read one step past the end of a task
  L.bidx.alloc(static_cast<size_t>(n_band), s, kPad);
  if (w) L.bw.alloc(static_cast<size_t>(n_band) * sizeof(RT), s, kPad * sizeof(RT));
  L.ppos.alloc(static_cast<size_t>(n_part), s);
  dbuf<uint8_t> hf8(static_cast<size_t>(n_band) + 8, s);
  dbuf<uint32_t> rank(static_cast<size_t>(n_band) + 1, s), fpos(static_cast<size_t>(n_part), s);
  fill_band_major_kernel<RT><<<bgrid(static_cast<size_t>(n_band)), kBT, 0, s>>>(
    keys, n_band, idx, m8.data(), w, bits, L.bidx.data(),
    w ? reinterpret_cast<RT*>(L.bw.data()) : nullptr, hf8.data());
  CUDA_TRY(cudaGetLastError());
  exclusive_scan<u8_one, uint32_t>(reinterpret_cast<const u8_one*>(hf8.data()), rank.data(),
                                   static_cast<size_t>(n_band), s, true);
  fill_ppos_kernel<<<bgrid(static_cast<size_t>(n_band)), kBT, 0, s>>>(
    keys, hf8.data(), rank.data(), pid.data(), n_band, L.ppos.data(), fpos.data());
  dbuf<int64_t> d_bstart(static_cast<size_t>(nb) + 1, s);
  band_starts_kernel<<<bgrid(static_cast<size_t>(nb) + 1), kBT, 0, s>>>(keys, n_band, nb,
                                                                        d_bstart.data());
  CUDA_TRY(cudaGetLastError());
  std::vector<int64_t> bstart(static_cast<size_t>(nb) + 1);
  uint32_t rank_total = 0;
  CUDA_TRY(cudaMemcpyAsync(bstart.data(), d_bstart.data(), bstart.size() * 8,
                           cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&rank_total, rank.data() + n_band, 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  B200_EXPECTS(static_cast<int64_t>(rank_total) == n_part, kUnknownError,
               "banded layout: cell counts of the two orders differ");

  // ---- tasks, groups, CTA ranges (host: a few hundred thousand windows) --------------------
  std::vector<int64_t> win_lo, win_hi;
  std::vector<int32_t> win_band;
  for (int k = 0; k < nb; ++k) {
    for (int64_t p = bstart[k]; p < bstart[k + 1]; p += kTaskWindow) {
      win_lo.push_back(p);
      win_hi.push_back(std::min<int64_t>(p + kTaskWindow, bstart[k + 1]));
      win_band.push_back(k);
    }
  }
  const int64_t ntasks = static_cast<int64_t>(win_lo.size());
  B200_EXPECTS(ntasks > 0 && ntasks < (int64_t{1} << 31), kUnknownError, "banded layout: no tasks");
  L.num_tasks = ntasks;
  {
    dbuf<int64_t> d_lo(win_lo.size(), s), d_hi(win_hi.size(), s);
    dbuf<int32_t> d_band(win_band.size(), s);
    CUDA_TRY(cudaMemcpyAsync(d_lo.data(), win_lo.data(), win_lo.size() * 8, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_hi.data(), win_hi.data(), win_hi.size() * 8, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(d_band.data(), win_band.data(), win_band.size() * 4,
                             cudaMemcpyHostToDevice, s));
    L.tasks.alloc(static_cast<size_t>(ntasks), s);
    finalize_tasks_kernel<<<bgrid(static_cast<size_t>(ntasks)), kBT, 0, s>>>(
      d_lo.data(), d_hi.data(), d_band.data(), ntasks, rank.data(), fpos.data(), n_band, n_part,
      L.tasks.data());
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));
  }
  // the work queue: units of consecutive tasks of one band in band order (a CTA that claims
  // consecutive units keeps its band in shared memory); unit sizes follow guided
  // self-scheduling, remaining / (2 * CTAs) clamped to [kUnitMin, kUnitMax] tasks, so that the
  // last units are small and the CTAs finish together
  const int nctas = h->sm_count;
  constexpr int64_t kUnitMin = 32, kUnitMax = 512;
  std::vector<band_group> groups;
  for (int64_t k = 0; k < ntasks;) {
    int64_t size = std::max<int64_t>(kUnitMin, std::min<int64_t>(kUnitMax, (ntasks - k) / (2 * nctas)));
    int64_t e    = std::min<int64_t>(ntasks, k + size);
    for (int64_t j = k + 1; j < e; ++j)
      if (win_band[j] != win_band[k]) { e = j; break; }
    band_group gpp{};
    gpp.task_first = static_cast<int32_t>(k);
    gpp.task_end   = static_cast<int32_t>(e);
    gpp.band       = win_band[k];
    groups.push_back(gpp);
    k = e;
  }
  L.num_groups = static_cast<int64_t>(groups.size());
  L.num_ctas   = nctas;
  L.groups.alloc(groups.size(), s);
  CUDA_TRY(cudaMemcpyAsync(L.groups.data(), groups.data(), groups.size() * sizeof(band_group),
                           cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));

  // ---- phase-2 schedule -------------------------------------------------------------------
  if (L.generic_rows) build_schedule(h, rp, tile_edges, row_cap);
  else build_schedule_reduced(h, rp, ep.seg[1], ep.seg[3], tile_edges, row_cap);
  L.enabled  = true;
  L.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin)
                 .count();
  if (std::getenv("B200_BAND_VERBOSE")) {
    fprintf(stderr,
            "[b200] banded layout: E=%lld banded=%lld (%.3f) partials=%lld reduced E'=%lld (%.3f) "
            "bands=%d tasks=%lld groups=%lld tiles=%lld chunks=%lld build=%.1f ms\n",
            (long long)ne, (long long)n_band, double(n_band) / double(ne), (long long)n_part,
            (long long)n_red, double(n_red) / double(ne), nb, (long long)ntasks,
            (long long)L.num_groups, (long long)rp.num_tiles, (long long)rp.num_chunks, L.build_ms);
  }
}

template <typename RT>
banded_layout* banded_get(b200_handle* h, edge_partition& ep, const RT* weights, int tile_edges,
                          int row_cap)
{
  const int slot = sizeof(RT) == 4 ? 0 : 1;
  if (!ep.banded[slot]) {
    auto L = std::make_shared<banded_layout>();
    build_banded<RT>(h, ep, weights, *L, tile_edges, row_cap);
    ep.banded[slot] = L;
  } else if (ep.banded[slot]->enabled) {
    // the layout is cached per value width; the tile shape of phase 2 may differ between calls
    edge_partition& rp = *ep.banded[slot]->reduced;
    if (rp.sched_tile_edges != tile_edges || rp.sched_row_cap != row_cap) {
      if (ep.banded[slot]->generic_rows) build_schedule(h, rp, tile_edges, row_cap);
      else build_schedule_reduced(h, rp, ep.seg[1], ep.seg[3], tile_edges, row_cap);
    }
  }
  return ep.banded[slot].get();
}

template banded_layout* banded_get<float>(b200_handle*, edge_partition&, const float*, int, int);
template banded_layout* banded_get<double>(b200_handle*, edge_partition&, const double*, int, int);

}  // namespace b200
