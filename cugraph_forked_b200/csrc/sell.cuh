// "Lane owns a segment": the round-2 layout and kernels of the two-phase pull SpMV.
//
// Round 1 (banded.cuh) stored the pre-reduced edges band-major with a cell-start flag in every
// index and reduced them with a segmented warp scan (0.7 - 1.0 warp instructions per edge: the
// kernel was bound by that bookkeeping, not by the shared-memory gathers, which alone take 112 us
// for the 215 M pre-reduced edges of an R-MAT scale-24 graph, profiles/r2_notes.md), and ran the
// warp-specialised tile kernel over the reduced CSC (5 - 12 ps per entry: staging, product
// write-back and row reductions through shared memory).  Here both phases are sliced-ELL
// kernels in which ONE LANE walks ONE segment with no flags, ballots, scans or scratch:
//
//   phase 1  a segment is a CELL (the edges of one row whose sources fall into one band of x).
//            The cells of a band are sorted by length (stable, so equal lengths stay in row
//            order) and packed 32 to a SLICE; the 16-bit band-local indices of a slice are
//            interleaved so that lane l reads its cell's next kSellG indices with one coalesced
//            8-byte load, gathers from the band's copy in shared memory and adds.  Without a
//            flag bit an index has all 16 bits, and without scratch the band gets the whole
//            shared memory: 57 344 fp32 / 28 672 fp64 entries (224 KB) instead of 32 K / 16 K,
//            and cells of 2 edges are worth pre-reducing (round 1: >= 3).  Padding slots point
//            at a zero entry behind the band, so the inner loop has no predicate.
//   phase 2  a segment is a ROW of the reduced CSC (its remaining direct sources + one entry per
//            pre-reduced cell).  Rows are sorted by reduced length inside windows of 1024 rows
//            (so a slice's 32 rows have nearly equal lengths and the epilogue's accesses stay
//            inside a 4 - 8 KB window), lane l gathers x[idx] / pval[idx - x_split] for its row
//            and runs the fused epilogue.  Rows with more than kSellLong entries are cut into
//            chunks (slices of their own) whose sums the tail kernel combines in a fixed order.
//
// The reference has no counterpart (per_v_transform_reduce_incoming_outgoing_e.cuh:194-479 walks
// the CSC rows directly).  Every row sum is the same sum in another FIXED association order (cells,
// slices, lanes and chunk order are fixed at build time; only the warp that runs a task is
// dynamic), so scores are reproducible run to run.  The two iteration scalars (L1 difference,
// dangling mass) are fp64 sums of per-CTA partials whose composition follows the dynamic task
// claims: they can differ in the last bit between runs (never seen to change an fp32 result or
// an iteration count; tests/test_gpu_centrality.py asserts bit-equal repeated runs).
#pragma once

#include "banded.cuh"
#include "ptx.cuh"

namespace b200 {

constexpr int kSellG        = 4;      // 16-bit indices per lane and load in phase 1
constexpr int kSellCellMax  = 512;    // largest cell (cells are cut at CSC positions that are multiples of the
                                      // build's cell_max: 512 for large graphs, 128 otherwise; sell.cu)
constexpr int kSellMinCell  = 2;      // cells with fewer edges stay direct gathers
constexpr int kSellLong     = 128;    // reduced rows with more entries are chunked
constexpr int kSellWindow   = 1024;   // rows per sort window of phase 2
constexpr int kSell2Threads = 256;
#ifndef B200_SELL2_MIN_CTAS
#define B200_SELL2_MIN_CTAS 4
#endif
constexpr int kSell2MinCtas = B200_SELL2_MIN_CTAS;   // CTAs per SM the phase-2 kernel is compiled for
constexpr int kSellQ1       = 32;     // largest phase-1 task quantum: load groups (of 32 * kSellG indices)
constexpr int kSellQ2       = 96;     // largest phase-2 task quantum: cost units (32 entries or 1/2 slice);
                                      // a task then has < 64 slices (sell_rows_kernel relies on it)

template <typename RT>
struct sell_geom {
  static constexpr int entries = sizeof(RT) == 4 ? 57344 : 28672;      // 224 KB
  static constexpr int pad     = entries;                              // index of the zero entry
  static constexpr size_t smem = (static_cast<size_t>(entries) + 16 / sizeof(RT)) * sizeof(RT);
};

struct sell_task {
  int32_t slice_first;
  int32_t slice_end;
  uint32_t g_first;   // first load group of slice_first
  int32_t n;          // load groups of EVERY slice of the task when they are all equal, else 0
};

struct sell_layout {
  int value_size{0};
  bool enabled{false};
  int band_entries{0};
  int32_t x_split{0};              // phase-2 indices >= x_split refer to partial (index - x_split)
  int64_t num_partials{0};         // pre-reduced cells; pval has one more (zero) entry
  int64_t num_banded_edges{0};
  bool skips_empty_rows{false};    // rows of all-empty slices are not stored: the caller clears the output
  bool band_major_slots{false};    // partial slots follow the sorted cell order (else row-major)
  // phase 1
  int64_t p1_slices{0};
  int64_t p1_groups{0};            // stored indices = p1_groups * 32 * kSellG
  dbuf<uint16_t> p1_idx;
  dbuf<char> p1_w;                 // RT, weighted graphs only
  dbuf<uint32_t> p1_sbase;         // [p1_slices + 1] first load group of every slice
  dbuf<int32_t> p1_slot;           // [p1_slices * 32] partial slot of every lane, -1 = none
  dbuf<sell_task> p1_tasks;
  dbuf<band_group> p1_units;       // work queue: tasks [task_first, task_end) of one band
  int64_t p1_num_tasks{0};
  int64_t p1_num_units{0};
  std::vector<int32_t> p1_unit_band;   // host copy: band of every unit (units are in band order)
  // phase 2
  int64_t p2_slices{0};
  int64_t p2_chunk_slices{0};      // slices [0, p2_chunk_slices) hold chunks of long rows
  int64_t p2_entries{0};           // stored entries = 32 * sum of slice lengths
  int64_t num_chunks{0};
  int64_t num_long_rows{0};
  int64_t direct_entries{0};
  dbuf<int32_t> p2_idx;
  dbuf<char> p2_w;
  dbuf<uint32_t> p2_sbase;         // [p2_slices + 1] in units of 32 entries
  dbuf<int32_t> p2_row;            // [p2_slices * 32] row of every lane, -1 = none
  dbuf<int32_t> p2_task_first;     // [p2_num_tasks + 1]
  int64_t p2_num_tasks{0};
  dbuf<int32_t> long_row_ids;      // ascending
  dbuf<int64_t> long_row_chunk_first;
  double build_ms{0.0};
};

template <typename RT>
struct sell_params1 {
  const RT* x;
  RT* pval;
  const uint16_t* idx;
  const RT* w;
  const uint32_t* sbase;
  const int32_t* slot;
  const sell_task* tasks;
  const band_group* units;
  int32_t unit_first;   // this launch runs the units [unit_first, unit_end)
  int32_t unit_end;
  int32_t* queue;
  int32_t* queue_next;
  int64_t num_minors;
  const iter_state* state;
};

template <typename RT>
struct sell_params2 {
  const RT* x;
  const RT* pval;
  int32_t x_split;
  const int32_t* idx;
  const RT* w;
  const uint32_t* sbase;
  const int32_t* row;
  int64_t chunk_slices;
  int64_t num_chunks;
  const int32_t* task_first;
  int32_t num_tasks;
  int32_t* queue;
  int32_t* queue_next;
  double* chunk_partials;
  double* cta_partials;
  int32_t slots;
  const iter_state* state;
  int32_t hot_limit;   // entries [0, hot_limit) of x are kept in L2 (evict_last), the rest streams
  int32_t use_hints;   // 0: plain read-only loads and stores (working sets that nearly fit L2)
};

// ------------------------------------------------------------------------------------------
// phase 1: one persistent CTA of 1024 threads per SM.  A CTA claims the next unit of the work
// queue (tasks of one band), loads the band of x into shared memory unless it already holds it,
// and its warps claim the unit's tasks (runs of slices).  Lane l of a warp walks cell l of the
// slice: per load group one 8-byte load of kSellG indices (coalesced: 256 B per warp), kSellG
// gathers from the band, kSellG independent accumulators; one store per cell at the end.
// ------------------------------------------------------------------------------------------
// slices [t.slice_first, t.slice_end) all have N load groups: U slices per step, all their index
// and slot loads issued before the first gather
template <typename RT, int N, int U>
__device__ __forceinline__ void sell_p1_uniform(sell_params1<RT> const& P, const RT* band,
                                                sell_task const& t, int lane)
{
  const uint2* ip0 = reinterpret_cast<const uint2*>(P.idx) + static_cast<int64_t>(t.g_first) * 32 + lane;
  for (int s = t.slice_first; s < t.slice_end; s += U) {
    uint2 c[U][N];
    int sl[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool live = s + u < t.slice_end;
      sl[u] = live ? __ldg(P.slot + static_cast<int64_t>(s + u) * 32 + lane) : -1;
#pragma unroll
      for (int g = 0; g < N; ++g)
        c[u][g] = live ? __ldcs(ip0 + (static_cast<int64_t>(s + u - t.slice_first) * N + g) * 32)
                       : make_uint2(0u, 0u);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      RT a0 = RT(0), a1 = RT(0), a2 = RT(0), a3 = RT(0);
#pragma unroll
      for (int g = 0; g < N; ++g) {
        a0 += band[c[u][g].x & 0xffffu];
        a1 += band[c[u][g].x >> 16];
        a2 += band[c[u][g].y & 0xffffu];
        a3 += band[c[u][g].y >> 16];
      }
      if (sl[u] >= 0) P.pval[sl[u]] = (a0 + a1) + (a2 + a3);
    }
  }
}

template <typename RT, bool WEIGHTED>
__global__ void __launch_bounds__(kBandThreads, 1) sell_band_kernel(const __grid_constant__ sell_params1<RT> P)
{
  using G = sell_geom<RT>;
  static_assert(kSellG == 4, "the index loads below are written for 4 x 16 bit");
  extern __shared__ __align__(16) unsigned char sell_smem[];
  RT* band = reinterpret_cast<RT*>(sell_smem);
  __shared__ int s_next;
  __shared__ int s_unit;
  __shared__ __align__(8) uint64_t s_bar;   // completion of the band's bulk copies

  const int tid = threadIdx.x, lane = tid & 31;
  if (blockIdx.x == 0 && tid == 0) *P.queue_next = 0;
  if (P.state && P.state->done) return;
  const RT* x = P.x;
  if (tid < static_cast<int>(16 / sizeof(RT))) band[G::entries + tid] = RT(0);   // the pad entry
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    fence_mbar_init();
  }
  uint32_t bar_parity = 0;

  int cur_band = -1;
  for (;;) {
    __syncthreads();  // every warp is done with the previous unit
    if (tid == 0) s_unit = atomicAdd(P.queue, 1);
    __syncthreads();
    const int u = P.unit_first + s_unit;
    if (u >= P.unit_end) break;
    const band_group grp = P.units[u];
    if (grp.band != cur_band) {
      cur_band            = grp.band;
      const int64_t first = static_cast<int64_t>(grp.band) * G::entries;
      const int64_t left  = P.num_minors - first;
      const int n         = static_cast<int>(left < G::entries ? left : G::entries);
      const RT* src       = x + first;
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        // the band comes in by bulk async copies (TMA 1D, SASS UBLKCP): one elected thread issues
        // them, the copy engine fills shared memory without a load instruction per 16 bytes, and
        // everybody waits on the mbarrier.  The generic-proxy reads of the previous band were
        // ordered by the __syncthreads above; the proxy fence orders them before these writes.
        constexpr int per     = 16 / sizeof(RT);
        const int nv          = n / per;                   // whole 16-byte units
        const uint32_t bytes  = static_cast<uint32_t>(nv) * 16u;
        if (tid == 0 && bytes) {
          fence_proxy_async();
          mbar_arrive_expect_tx(&s_bar, bytes);
          const uint64_t pol = policy_evict_last();        // other CTAs load the same band
          constexpr uint32_t kChunk = 32768;
          for (uint32_t o = 0; o < bytes; o += kChunk)
            bulk_g2s(reinterpret_cast<char*>(band) + o, reinterpret_cast<const char*>(src) + o,
                     bytes - o < kChunk ? bytes - o : kChunk, &s_bar, pol);
        }
        for (int i = nv * per + tid; i < n; i += kBandThreads) band[i] = src[i];
        if (bytes) mbar_wait(&s_bar, bar_parity);
        if (bytes) bar_parity ^= 1u;
      } else {
        for (int i = tid; i < n; i += kBandThreads) band[i] = src[i];
      }
    }
    if (tid == 0) s_next = grp.task_first;
    __syncthreads();
    for (;;) {
      int ti = 0;
      if (lane == 0) ti = atomicAdd(&s_next, 1);
      ti = __shfl_sync(0xffffffffu, ti, 0);
      if (ti >= grp.task_end) break;
      const sell_task t = P.tasks[ti];
      if constexpr (!WEIGHTED) {
        // most slices hold short cells (1 - 2 load groups): tasks whose slices all have the same
        // number of groups need no per-slice look-up, and several slices go in flight at once
        if (t.n == 1) { sell_p1_uniform<RT, 1, 4>(P, band, t, lane); continue; }
        if (t.n == 2) { sell_p1_uniform<RT, 2, 2>(P, band, t, lane); continue; }
        if (t.n == 3) { sell_p1_uniform<RT, 3, 2>(P, band, t, lane); continue; }
        if (t.n == 4) { sell_p1_uniform<RT, 4, 2>(P, band, t, lane); continue; }
      }
      for (int s = t.slice_first; s < t.slice_end; ++s) {
        const uint32_t g0 = __ldg(P.sbase + s), g1 = __ldg(P.sbase + s + 1);
        const int sl      = __ldg(P.slot + static_cast<int64_t>(s) * 32 + lane);
        const uint2* ip   = reinterpret_cast<const uint2*>(P.idx) + static_cast<int64_t>(g0) * 32 + lane;
        const int n       = static_cast<int>(g1 - g0);
        RT a0 = RT(0), a1 = RT(0), a2 = RT(0), a3 = RT(0);
        if constexpr (!WEIGHTED) {
#pragma unroll 4
          for (int g = 0; g < n; ++g) {
            const uint2 c = __ldcs(ip + static_cast<int64_t>(g) * 32);
            a0 += band[c.x & 0xffffu];
            a1 += band[c.x >> 16];
            a2 += band[c.y & 0xffffu];
            a3 += band[c.y >> 16];
          }
        } else {
          const RT* wp = P.w + (static_cast<int64_t>(g0) * 32 + lane) * kSellG;
#pragma unroll 2
          for (int g = 0; g < n; ++g) {
            const uint2 c = __ldcs(ip + static_cast<int64_t>(g) * 32);
            RT w0, w1, w2, w3;
            if constexpr (sizeof(RT) == 4) {
              const float4 w = __ldcs(reinterpret_cast<const float4*>(wp + static_cast<int64_t>(g) * 32 * kSellG));
              w0 = w.x; w1 = w.y; w2 = w.z; w3 = w.w;
            } else {
              const double2* w2p = reinterpret_cast<const double2*>(wp + static_cast<int64_t>(g) * 32 * kSellG);
              const double2 wa = __ldcs(w2p), wb = __ldcs(w2p + 1);
              w0 = wa.x; w1 = wa.y; w2 = wb.x; w3 = wb.y;
            }
            a0 += band[c.x & 0xffffu] * w0;
            a1 += band[c.x >> 16] * w1;
            a2 += band[c.y & 0xffffu] * w2;
            a3 += band[c.y >> 16] * w3;
          }
        }
        if (sl >= 0) P.pval[sl] = (a0 + a1) + (a2 + a3);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// phase 2: 256-thread CTAs, no shared memory (everything goes to L1), warps claim tasks (runs
// of slices) from a global counter.  Lane l walks row l of the slice: one coalesced 4-byte
// index load per entry (128 B per warp), one gather from x or from the partial sums, two
// accumulators; then the fused epilogue of the row (inputs prefetched before the walk).
// ------------------------------------------------------------------------------------------
template <typename RT, bool WEIGHTED, typename EPI>
__global__ void __launch_bounds__(kSell2Threads, kSell2MinCtas) sell_rows_kernel(
  const __grid_constant__ sell_params2<RT> P, EPI epi)
{
  __shared__ double red_scratch[32];
  const int tid = threadIdx.x, lane = tid & 31;
  if (blockIdx.x == 0 && tid == 0) *P.queue_next = 0;
  if (P.state && P.state->done) return;
  epi.setup(P.state);
  acc2 acc;
  const RT* x  = P.x;
  const RT* pv = P.pval - P.x_split;   // indexed by the entry itself
  // L2 replacement order: the head of x (degree-sorted: the sources most rows gather) stays, its
  // cold tail streams through, the partial sums (eight rows to a line) are ordinary lines
  const uint64_t pol_hot = policy_evict_last(), pol_cold = policy_evict_first(),
                 pol_part = policy_evict_normal();
  const bool hints = P.use_hints != 0;   // uniform over the grid
  auto gather = [&](int32_t i) -> RT {
    const bool part = i >= P.x_split;
    const RT* a     = (part ? pv : x) + i;
    return hints ? ld_policy(a, part ? pol_part : (i < P.hot_limit ? pol_hot : pol_cold)) : __ldg(a);
  };
  for (;;) {
    int ti = 0;
    if (lane == 0) ti = atomicAdd(P.queue, 1);
    ti = __shfl_sync(0xffffffffu, ti, 0);
    if (ti >= P.num_tasks) break;
    const int s_first = __ldg(P.task_first + ti), s_end = __ldg(P.task_first + ti + 1);
    const int ns      = s_end - s_first;   // <= kSellQ2 / 2 < 64 slices (every slice costs >= 2)
    // the slice bounds of the whole task in two registers per lane: no dependent look-up later
    const uint32_t sb_lo = lane <= ns ? __ldg(P.sbase + s_first + lane) : 0u;
    const uint32_t sb_hi = 32 + lane <= ns ? __ldg(P.sbase + s_first + 32 + lane) : 0u;
    auto bound = [&](int i) {
      const uint32_t lo = __shfl_sync(0xffffffffu, sb_lo, i & 31);
      const uint32_t hi = __shfl_sync(0xffffffffu, sb_hi, i & 31);
      return i < 32 ? lo : hi;
    };
    // two slices per step: their row ids and first entries are all loaded before anything that
    // depends on them
    for (int k = 0; k < ns; k += 2) {
      const bool two = k + 1 < ns;
      uint32_t b[3];
      b[0] = bound(k);
      b[1] = bound(k + 1);
      b[2] = two ? bound(k + 2) : b[1];
      int row[2];
      bool chunk[2];
      int32_t e[2][4];
      const int32_t* ip[2];
      int n[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int sidx = s_first + k + u;
        n[u]     = static_cast<int>(b[u + 1] - b[u]);
        chunk[u] = sidx < P.chunk_slices;
        const bool live = u == 0 || two;
        row[u]   = (chunk[u] || !live) ? -1 : __ldcs(P.row + static_cast<int64_t>(sidx) * 32 + lane);
        ip[u]    = P.idx + static_cast<int64_t>(b[u]) * 32 + lane;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          e[u][j] = j < n[u] ? __ldcs(ip[u] + j * 32) : -1;
      }
      typename EPI::pre_t pre[2];
      RT a0[2], a1[2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        pre[u] = typename EPI::pre_t{};
        if (row[u] >= 0) pre[u] = hints ? epi.prefetch_stream(row[u]) : epi.prefetch(row[u]);
        RT v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = RT(0);
          if (e[u][j] >= 0) v[j] = gather(e[u][j]);
        }
        if constexpr (WEIGHTED) {
          const RT* wp = P.w + static_cast<int64_t>(b[u]) * 32 + lane;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j < n[u]) v[j] *= __ldcs(wp + j * 32);
        }
        a0[u] = v[0] + v[2];
        a1[u] = v[1] + v[3];
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        // the rest of a longer slice
        int j = 4;
        for (; j + 4 <= n[u]; j += 4) {
          const int32_t i0 = __ldcs(ip[u] + static_cast<int64_t>(j) * 32);
          const int32_t i1 = __ldcs(ip[u] + static_cast<int64_t>(j + 1) * 32);
          const int32_t i2 = __ldcs(ip[u] + static_cast<int64_t>(j + 2) * 32);
          const int32_t i3 = __ldcs(ip[u] + static_cast<int64_t>(j + 3) * 32);
          RT v0 = gather(i0), v1 = gather(i1), v2 = gather(i2), v3 = gather(i3);
          if constexpr (WEIGHTED) {
            const RT* wp = P.w + static_cast<int64_t>(b[u]) * 32 + lane;
            v0 *= __ldcs(wp + static_cast<int64_t>(j) * 32);
            v1 *= __ldcs(wp + static_cast<int64_t>(j + 1) * 32);
            v2 *= __ldcs(wp + static_cast<int64_t>(j + 2) * 32);
            v3 *= __ldcs(wp + static_cast<int64_t>(j + 3) * 32);
          }
          a0[u] += v0; a1[u] += v1; a0[u] += v2; a1[u] += v3;
        }
        for (; j < n[u]; ++j) {
          const int32_t i0 = __ldcs(ip[u] + static_cast<int64_t>(j) * 32);
          RT v0 = gather(i0);
          if constexpr (WEIGHTED)
            v0 *= __ldcs(P.w + static_cast<int64_t>(b[u]) * 32 + lane + static_cast<int64_t>(j) * 32);
          a0[u] += v0;
        }
        const RT sum = a0[u] + a1[u];
        if (chunk[u]) {
          const int64_t c = static_cast<int64_t>(s_first + k + u) * 32 + lane;
          if ((u == 0 || two) && c < P.num_chunks) P.chunk_partials[c] = static_cast<double>(sum);
        } else if (row[u] >= 0) {
          if (hints) epi.finish_stream(row[u], sum, pre[u], acc, P.hot_limit);
          else epi.finish(row[u], sum, pre[u], acc);
        }
      }
    }
  }
  if (!P.cta_partials) return;
  const double a = block_sum(acc.a, red_scratch);
  const double b = block_sum(acc.b, red_scratch);
  if (tid == 0) {
    P.cta_partials[blockIdx.x]           = a;
    P.cta_partials[P.slots + blockIdx.x] = b;
  }
}

// builds (or returns the cached) layout for value width sizeof(RT); sell.cu
template <typename RT>
sell_layout* sell_get(b200_handle* h, edge_partition& ep, const RT* weights, bool skip_empty = false);

}  // namespace b200
