// Banded pre-reduction: phase 1 of the two-phase pull SpMV (single GPU).
//
// Why.  The tiled pull kernel (spmv.cuh) is bound by the L2 -> SM sector path of its random
// 4/8-byte gathers (one 32-byte sector per gathered value, ~1 sector/clk/SM),
// not by HBM.  A B200 has 148 x 227 KB of shared memory - more than a third of the gathered
// vector of a scale-24 graph - and random gathers from shared memory run several times faster
// than through L1TEX.  So the gathered vector x is cut into BANDS of 2^band_bits entries
// (128 KB), and every row's ascending neighbour list into CELLS = maximal runs of sources
// inside one band.  Cells with >= kMinCell edges (80 % of the edges of an R-MAT scale-24 graph,
// in 0.05 cells per edge: multi-edges and hub sources) are PRE-REDUCED: their edges are stored
// band-major as 16-bit band-local indices (bit 15 = "first edge of a cell"), a CTA loads a band
// of x into shared memory once and streams the band's edges through it with coalesced 128-bit
// loads, and writes one partial sum per cell.  Phase 2 is the unchanged tiled pull kernel over
// a REDUCED CSC in which every pre-reduced cell is a single entry that refers to its partial
// sum (index >= x_split reads the partial vector); the remaining cells keep their edges.
// Phase 2 therefore gathers 0.25 values per edge instead of 1.0, and a row's partials are
// consecutive in memory (eight to a sector).
//
// The reference has no counterpart: its kernels gather x[minor] per edge from global memory
// (cpp/src/prims/per_v_transform_reduce_incoming_outgoing_e.cuh:194-400).  Results are the same
// sums in a different (fixed) association order; everything is deterministic: cells, tasks and
// the lane mapping are fixed at build time, only the warp that runs a task is dynamic.
#pragma once

#include "common.cuh"
#include "epilogues.cuh"
#include "prims.cuh"

namespace b200 {

constexpr int kBandThreads = 1024;
constexpr int kBandWarps   = kBandThreads / 32;
#ifndef B200_MIN_CELL
#define B200_MIN_CELL 3
#endif
constexpr int kMinCell     = B200_MIN_CELL;  // cells with fewer edges stay direct gathers
constexpr int kLaneFlags   = (8 + kMinCell - 1) / kMinCell;  // most cell starts in 8 edges
constexpr int kCellMax     = 2048;   // cells are cut at CSC positions that are multiples of this
#ifndef B200_TASK_WINDOW
#define B200_TASK_WINDOW 4096
#endif
constexpr int kTaskWindow  = B200_TASK_WINDOW;  // a task = the cells that start inside one window of a band
static_assert(kTaskWindow >= kCellMax, "every window must contain a cell start");
constexpr int kStepEdges   = 256;    // edges a warp handles per step (8 per lane, one LDG.128)
constexpr int64_t kBandLoadCost = 16384;  // cost of loading a band, in edges (CTA balancing)
constexpr int64_t kHotBandMin   = 32768;  // bands with fewer pre-reducible edges stay direct

template <typename RT>
struct band_geom {
  static constexpr int bits    = sizeof(RT) == 4 ? 15 : 14;   // 128 KB of shared memory
  static constexpr int entries = 1 << bits;
  static constexpr uint32_t mask = entries - 1;
};

// one unit of work of a warp: a run of whole cells of one band
struct alignas(8) band_task {
  int64_t first;        // position of the first edge in the band-major arrays (a cell start)
  int64_t rank_first;   // band-major rank of that cell (index into ppos)
  int32_t count;        // edges
  int32_t band;
};
static_assert(sizeof(band_task) == 24, "band_task layout");

// one unit of the work queue: tasks [task_first, task_end) of one band, run by one CTA with the
// band in shared memory.  CTAs claim units in order with a global counter; unit sizes shrink
// towards the end of the queue (guided self-scheduling).
struct band_group {
  int32_t task_first;
  int32_t task_end;
  int32_t band;
  int32_t pad;
};

struct banded_layout {
  int band_bits{0};
  int value_size{0};
  bool enabled{false};          // false: the graph has too little to pre-reduce
  // rows of a multi-GPU partition follow the GLOBAL degree order: the reduced partition is then
  // scheduled by the generic host pass instead of build_schedule_reduced
  bool generic_rows{false};
  int64_t num_banded_edges{0};
  int64_t num_partials{0};
  int64_t num_tasks{0};
  int64_t num_groups{0};
  int num_ctas{0};
  int32_t x_split{0};           // reduced indices >= x_split refer to partial (index - x_split)
  dbuf<uint16_t> bidx;          // [num_banded_edges + pad] band-local index | cell-start << 15
  dbuf<char> bw;                // RT [num_banded_edges + pad], weighted graphs only
  dbuf<int32_t> ppos;           // [num_partials] band-major cell rank -> partial slot (row-major)
  dbuf<band_task> tasks;
  dbuf<band_group> groups;        // the work queue: units of consecutive tasks of one band
  std::unique_ptr<edge_partition> reduced;  // phase-2 CSC + its tile schedule
  double build_ms{0.0};
};

struct band_params {
  const void* x;                  // RT [num_minors]
  void* pval;                     // RT [num_partials]
  const uint16_t* bidx;
  const void* bw;                 // RT or nullptr
  const int32_t* ppos;
  const band_task* tasks;
  const band_group* groups;
  int32_t num_groups;
  int32_t* queue;                 // this launch's unit counter (0 at launch)
  int32_t* queue_next;            // the next launch's counter, reset by this launch
  int64_t num_minors;
  const iter_state* state;        // may be nullptr
};

// ------------------------------------------------------------------------------------------
// phase-1 kernel: one persistent CTA per SM.  A CTA claims the next unit of the work queue,
// loads the unit's band of x into shared memory unless it already holds it (units are queued
// in band order), and its 32 warps then claim the unit's tasks.  A warp walks a task in
// steps of 256 edges (lane l owns edges [8l, 8l + 8) of the step): one 128-bit load of eight
// 16-bit indices (prefetched a step ahead), eight shared-memory gathers, a sequential
// segmented sum inside the lane, a segmented warp scan of the lanes' open tails, and one store
// per cell that ends in the lane.  The sum that is still open at the end of a step is carried
// in a register to the next step.
// ------------------------------------------------------------------------------------------
// flag bits (bit 15 of every 16-bit index) of the eight indices of a 128-bit load -> bits 0..7
__device__ __forceinline__ uint32_t band_flag_bits(uint4 const& c)
{
  // bytes 1 and 3 of every word carry the flags in their top bit
  const uint32_t lo = __byte_perm(c.x, c.y, 0x7531);  // [x.b1, x.b3, y.b1, y.b3]
  const uint32_t hi = __byte_perm(c.z, c.w, 0x7531);
  const uint32_t a  = ((lo >> 7) & 0x01010101u) * 0x01020408u;  // bits 24..27 = the four flags
  const uint32_t b  = ((hi >> 7) & 0x01010101u) * 0x01020408u;
  return (a >> 24) | ((b >> 24) << 4);
}

template <typename RT, bool WEIGHTED>
__global__ void __launch_bounds__(kBandThreads, 1) band_reduce_kernel(const __grid_constant__ band_params P)
{
  using G = band_geom<RT>;
  static_assert(kMinCell >= 2 && kLaneFlags <= 4, "a lane of 8 edges holds at most 4 cell starts");
  extern __shared__ __align__(16) unsigned char band_smem[];
  RT* band = reinterpret_cast<RT*>(band_smem);
  __shared__ int s_next;
  __shared__ int s_unit;

  const int tid = threadIdx.x, lane = tid & 31;
  // the other queue counter is the next launch's: reset it (stream order makes this safe)
  if (blockIdx.x == 0 && tid == 0) *P.queue_next = 0;
  if (P.state && P.state->done) return;
  const uint32_t lanes_lt = (1u << lane) - 1u;
  const RT* x   = static_cast<const RT*>(P.x);
  RT* pval      = static_cast<RT*>(P.pval);
  const RT* bw  = static_cast<const RT*>(P.bw);

  int cur_band = -1;
  for (;;) {
    __syncthreads();  // every warp is done with the previous unit
    if (tid == 0) s_unit = atomicAdd(P.queue, 1);
    __syncthreads();
    const int u = s_unit;
    if (u >= P.num_groups) break;
    const band_group grp = P.groups[u];
    if (grp.band != cur_band) {
      cur_band            = grp.band;
      const int64_t first = static_cast<int64_t>(grp.band) << G::bits;
      const int64_t left  = P.num_minors - first;
      const int n         = static_cast<int>(left < G::entries ? left : G::entries);
      const RT* src       = x + first;
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        constexpr int per = 16 / sizeof(RT);
        const int nv      = n / per;
        const int4* s4    = reinterpret_cast<const int4*>(src);
        int4* d4          = reinterpret_cast<int4*>(band);
        for (int i = tid; i < nv; i += kBandThreads) d4[i] = __ldg(s4 + i);
        for (int i = nv * per + tid; i < n; i += kBandThreads) band[i] = src[i];
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
      const band_task t  = P.tasks[ti];
      const int64_t a0   = t.first & ~int64_t(7);          // 16-byte aligned start
      const int lead     = static_cast<int>(t.first - a0);
      const int total    = lead + t.count;
      const int steps    = (total + kStepEdges - 1) / kStepEdges;
      const uint16_t* ip = P.bidx + a0 + lane * 8;
      const int32_t* pp  = P.ppos + t.rank_first;   // slots of this task's cells
      int rank           = 0;              // cell starts met so far (relative to rank_first)
      RT carry           = RT(0);          // sum of the cell that is open at the step boundary
      uint4 cur          = __ldcs(reinterpret_cast<const uint4*>(ip));
      for (int st = 0; st < steps; ++st) {
        uint4 nxt = make_uint4(0, 0, 0, 0);
        if (st + 1 < steps) nxt = __ldcs(reinterpret_cast<const uint4*>(ip + (st + 1) * kStepEdges));
        const int base = st * kStepEdges + lane * 8;   // position of the lane's first edge
        // gathers from the band (every index is a valid band offset, also outside the task)
        RT v[8];
        v[0] = band[cur.x & G::mask]; v[1] = band[(cur.x >> 16) & G::mask];
        v[2] = band[cur.y & G::mask]; v[3] = band[(cur.y >> 16) & G::mask];
        v[4] = band[cur.z & G::mask]; v[5] = band[(cur.z >> 16) & G::mask];
        v[6] = band[cur.w & G::mask]; v[7] = band[(cur.w >> 16) & G::mask];
        uint32_t fmask = band_flag_bits(cur);
        if (WEIGHTED) {
          // the weight arrays are padded, so this never reads out of bounds
          const RT* wp = bw + a0 + base;
          if constexpr (sizeof(RT) == 4) {
            const float4 w0 = __ldcs(reinterpret_cast<const float4*>(wp));
            const float4 w1 = __ldcs(reinterpret_cast<const float4*>(wp) + 1);
            v[0] *= w0.x; v[1] *= w0.y; v[2] *= w0.z; v[3] *= w0.w;
            v[4] *= w1.x; v[5] *= w1.y; v[6] *= w1.z; v[7] *= w1.w;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const double2 w = __ldcs(reinterpret_cast<const double2*>(wp) + i);
              v[2 * i] *= w.x; v[2 * i + 1] *= w.y;
            }
          }
        }
        if (base < lead || base + 8 > total) {
          // first / last lanes of a task: mask what lies outside
          uint32_t vmask = 0;
#pragma unroll
          for (int i = 0; i < 8; ++i) vmask |= ((base + i >= lead && base + i < total) ? 1u : 0u) << i;
          fmask &= vmask;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (!(vmask & (1u << i))) v[i] = RT(0);
        }
        // ranks of the lane's cell starts: cells have >= kMinCell edges, so a lane holds at most
        // kLaneFlags of them and 2-3 ballots give the exclusive prefix of the counts
        const int cnt     = __popc(fmask);
        const uint32_t b0 = __ballot_sync(0xffffffffu, cnt & 1);
        const uint32_t b1 = __ballot_sync(0xffffffffu, cnt & 2);
        int excl          = __popc(b0 & lanes_lt) + 2 * __popc(b1 & lanes_lt);
        int step_flags    = __popc(b0) + 2 * __popc(b1);
        if constexpr (kLaneFlags > 3) {
          const uint32_t b2 = __ballot_sync(0xffffffffu, cnt & 4);
          excl += 4 * __popc(b2 & lanes_lt);
          step_flags += 4 * __popc(b2);
        }
        // slot of the cell that ENDS at the lane's k-th cell start (= the previous cell); the
        // task's very first cell start ends nothing.  Loaded now, used after the reduction.
        const int rb   = rank + excl - 1;
        const int skip = (rb < 0) ? 1 : 0;
        int sl[kLaneFlags];
#pragma unroll
        for (int k = 0; k < kLaneFlags; ++k) {
          sl[k] = 0;
          if (cnt > k + skip) sl[k] = __ldg(pp + rb + k + skip);   // k-th cell start that stores
        }
        // sequential segmented sum inside the lane; c[k] = the sum closed at the lane's k-th cell
        // start (cell starts are >= kMinCell apart, so start k sits at position >= k * kMinCell)
        RT c[kLaneFlags];
#pragma unroll
        for (int k = 0; k < kLaneFlags; ++k) c[k] = RT(0);
        RT run = RT(0);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (fmask & (1u << i)) {
            if (i < kMinCell) {
              c[0] = run;
            } else {
              const int nb = __popc(fmask & ((1u << i) - 1u));
#pragma unroll
              for (int k = 0; k < kLaneFlags; ++k)
                if (k * kMinCell <= i && nb == k) c[k] = run;
            }
            run = RT(0);
          }
          run += v[i];
        }
        // segmented inclusive scan of the open tails across the lanes: lane l adds the tails of
        // the lanes back to the nearest lane that holds a cell start
        const uint32_t fl = __ballot_sync(0xffffffffu, cnt > 0);
        const uint32_t le = fl & (lanes_lt | (1u << lane));
        const int seg0    = le ? 31 - __clz(le) : 0;   // first lane of this lane's segment
        RT sv = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const RT pv = __shfl_up_sync(0xffffffffu, sv, o);
          if (lane - o >= seg0) sv += pv;
        }
        // what is open when the lane's first edge is reached
        const RT up_v = __shfl_up_sync(0xffffffffu, sv, 1);
        RT carry_in   = carry;
        if (lane > 0) carry_in = (fl & lanes_lt) ? up_v : up_v + carry;
        const RT last_v = __shfl_sync(0xffffffffu, sv, 31);
        // stores: the lane's first cell start closes the cell carried into the lane, unless it
        // is the task's first cell start
        if (cnt > 0) {
          c[0] += carry_in;
          if (skip) {
#pragma unroll
            for (int k = 0; k + 1 < kLaneFlags; ++k) c[k] = c[k + 1];
          }
#pragma unroll
          for (int k = 0; k < kLaneFlags; ++k)
            if (cnt > k + skip) pval[sl[k]] = c[k];
        }
        carry = fl ? last_v : last_v + carry;
        rank += step_flags;
        cur = nxt;
      }
      // the cell that is still open at the end of the task
      if (lane == 0 && rank > 0) pval[__ldg(pp + rank - 1)] = carry;
    }
  }
}

// ------------------------------------------------------------------------------------------
// phase-1 kernel, second formulation (used for unweighted 4-byte values, where it measured 3 %
// faster; the first formulation is faster for the other types): LE = 16 (4-byte values) or 8
// (8-byte values) edges per lane and step, and the per-position running sums go through a
// private column of a shared-memory scratch array instead of registers, so that the cells that
// end in a lane are handled by a short loop over the lane's cell-start bits instead of
// predicated code for every position.  Twice the edges per step halve the per-edge cost of the
// ballots and of the segmented scan: 0.72 warp instructions per pre-reduced edge instead of
// 1.0 - but the kernel then runs into the shared-memory pipe (76 % busy: the random band
// gathers cost 5.3 wavefronts each), profiles/r1l_ncu_band_reduce2.md.
// ------------------------------------------------------------------------------------------
template <typename RT>
struct band_geom2 {
  static constexpr int LE         = sizeof(RT) == 4 ? 16 : 8;   // edges per lane and step
  static constexpr int step_edges = 32 * LE;
  static constexpr size_t smem    = size_t(band_geom<RT>::entries) * sizeof(RT) +
                                 size_t(LE) * kBandThreads * sizeof(RT);   // band + scratch
};

template <typename RT, bool WEIGHTED>
__global__ void __launch_bounds__(kBandThreads, 1) band_reduce2_kernel(const __grid_constant__ band_params P)
{
  using G  = band_geom<RT>;
  using G2 = band_geom2<RT>;
  constexpr int LE = G2::LE;
  constexpr int SE = G2::step_edges;
  static_assert((LE + kMinCell - 1) / kMinCell <= 7, "three ballots count at most 7 cell starts");
  extern __shared__ __align__(16) unsigned char band_smem[];
  RT* band = reinterpret_cast<RT*>(band_smem);
  __shared__ int s_next;
  __shared__ int s_unit;

  const int tid = threadIdx.x, lane = tid & 31;
  if (blockIdx.x == 0 && tid == 0) *P.queue_next = 0;
  if (P.state && P.state->done) return;
  RT* my = band + G::entries + tid;   // scratch column of this thread: my[i * kBandThreads]
  const uint32_t lanes_lt = (1u << lane) - 1u;
  const RT* x   = static_cast<const RT*>(P.x);
  RT* pval      = static_cast<RT*>(P.pval);
  const RT* bw  = static_cast<const RT*>(P.bw);

  int cur_band = -1;
  for (;;) {
    __syncthreads();  // every warp is done with the previous unit
    if (tid == 0) s_unit = atomicAdd(P.queue, 1);
    __syncthreads();
    const int u = s_unit;
    if (u >= P.num_groups) break;
    const band_group grp = P.groups[u];
    if (grp.band != cur_band) {
      cur_band            = grp.band;
      const int64_t first = static_cast<int64_t>(grp.band) << G::bits;
      const int64_t left  = P.num_minors - first;
      const int n         = static_cast<int>(left < G::entries ? left : G::entries);
      const RT* src       = x + first;
      if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        constexpr int per = 16 / sizeof(RT);
        const int nv      = n / per;
        const int4* s4    = reinterpret_cast<const int4*>(src);
        int4* d4          = reinterpret_cast<int4*>(band);
        for (int i = tid; i < nv; i += kBandThreads) d4[i] = __ldg(s4 + i);
        for (int i = nv * per + tid; i < n; i += kBandThreads) band[i] = src[i];
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
      const band_task t  = P.tasks[ti];
      const int64_t a0   = t.first & ~int64_t(7);          // 16-byte aligned start
      const int lead     = static_cast<int>(t.first - a0);
      const int total    = lead + t.count;
      const int steps    = (total + SE - 1) / SE;
      const uint16_t* ip = P.bidx + a0 + lane * LE;
      const int32_t* pp  = P.ppos + t.rank_first;   // slots of this task's cells
      int rank           = 0;              // cell starts met so far (relative to rank_first)
      RT carry           = RT(0);          // sum of the cell that is open at the step boundary
      uint4 cur[LE / 8], nxt[LE / 8];
#pragma unroll
      for (int q = 0; q < LE / 8; ++q) cur[q] = __ldcs(reinterpret_cast<const uint4*>(ip) + q);
      for (int st = 0; st < steps; ++st) {
#pragma unroll
        for (int q = 0; q < LE / 8; ++q) nxt[q] = make_uint4(0, 0, 0, 0);
        if (st + 1 < steps) {
#pragma unroll
          for (int q = 0; q < LE / 8; ++q)
            nxt[q] = __ldcs(reinterpret_cast<const uint4*>(ip + (st + 1) * SE) + q);
        }
        const int base = st * SE + lane * LE;   // position of the lane's first edge
        // gathers from the band (every index is a valid band offset, also outside the task)
        RT v[LE];
        uint32_t fmask = 0;
#pragma unroll
        for (int q = 0; q < LE / 8; ++q) {
          const uint4 c = cur[q];
          v[8 * q + 0] = band[c.x & G::mask]; v[8 * q + 1] = band[(c.x >> 16) & G::mask];
          v[8 * q + 2] = band[c.y & G::mask]; v[8 * q + 3] = band[(c.y >> 16) & G::mask];
          v[8 * q + 4] = band[c.z & G::mask]; v[8 * q + 5] = band[(c.z >> 16) & G::mask];
          v[8 * q + 6] = band[c.w & G::mask]; v[8 * q + 7] = band[(c.w >> 16) & G::mask];
          fmask |= band_flag_bits(c) << (8 * q);
        }
        if (WEIGHTED) {
          // the weight arrays are padded, so this never reads out of bounds
          const RT* wp = bw + a0 + base;
          constexpr int per = 16 / sizeof(RT);
#pragma unroll
          for (int q = 0; q < LE / per; ++q) {
            if constexpr (sizeof(RT) == 4) {
              const float4 w = __ldcs(reinterpret_cast<const float4*>(wp) + q);
              v[4 * q] *= w.x; v[4 * q + 1] *= w.y; v[4 * q + 2] *= w.z; v[4 * q + 3] *= w.w;
            } else {
              const double2 w = __ldcs(reinterpret_cast<const double2*>(wp) + q);
              v[2 * q] *= w.x; v[2 * q + 1] *= w.y;
            }
          }
        }
        if (base < lead || base + LE > total) {
          // first / last lanes of a task: mask what lies outside
          uint32_t vmask = 0;
#pragma unroll
          for (int i = 0; i < LE; ++i) vmask |= ((base + i >= lead && base + i < total) ? 1u : 0u) << i;
          fmask &= vmask;
#pragma unroll
          for (int i = 0; i < LE; ++i)
            if (!(vmask & (1u << i))) v[i] = RT(0);
        }
        // ranks of the lane's cell starts (<= 7 per lane): three ballots give the exclusive prefix
        const int cnt     = __popc(fmask);
        const uint32_t b0 = __ballot_sync(0xffffffffu, cnt & 1);
        const uint32_t b1 = __ballot_sync(0xffffffffu, cnt & 2);
        const uint32_t b2 = __ballot_sync(0xffffffffu, cnt & 4);
        const int excl = __popc(b0 & lanes_lt) + 2 * __popc(b1 & lanes_lt) + 4 * __popc(b2 & lanes_lt);
        const int step_flags = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
        // sequential segmented sum inside the lane; the running sum BEFORE every position goes to
        // the thread's scratch column (only the entries at cell starts are read back)
        RT run = RT(0);
#pragma unroll
        for (int i = 0; i < LE; ++i) {
          my[i * kBandThreads] = run;   // unconditional: measured faster than a predicated store
          if (fmask & (1u << i)) run = RT(0);
          run += v[i];
        }
        // segmented inclusive scan of the open tails across the lanes
        const uint32_t fl = __ballot_sync(0xffffffffu, cnt > 0);
        const uint32_t le = fl & (lanes_lt | (1u << lane));
        const int seg0    = le ? 31 - __clz(le) : 0;   // first lane of this lane's segment
        RT sv = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const RT pv = __shfl_up_sync(0xffffffffu, sv, o);
          if (lane - o >= seg0) sv += pv;
        }
        const RT up_v = __shfl_up_sync(0xffffffffu, sv, 1);
        RT carry_in   = carry;
        if (lane > 0) carry_in = (fl & lanes_lt) ? up_v : up_v + carry;
        const RT last_v = __shfl_sync(0xffffffffu, sv, 31);
        // the cells that end in this lane: one store per cell start, the first one closes the
        // cell carried into the lane (the task's very first cell start closes nothing)
        {
          int r      = rank + excl - 1;
          uint32_t m = fmask;
          RT add     = carry_in;
          while (m) {
            const int i = __ffs(m) - 1;
            m &= m - 1;
            const RT val = my[i * kBandThreads] + add;
            add          = RT(0);
            if (r >= 0) pval[__ldg(pp + r)] = val;
            ++r;
          }
        }
        carry = fl ? last_v : last_v + carry;
        rank += step_flags;
#pragma unroll
        for (int q = 0; q < LE / 8; ++q) cur[q] = nxt[q];
      }
      // the cell that is still open at the end of the task
      if (lane == 0 && rank > 0) pval[__ldg(pp + rank - 1)] = carry;
    }
  }
}

// builds (or returns the cached) layout for value width sizeof(RT); `weights` is the RT-typed
// weight array of the partition or nullptr.  banded.cu
template <typename RT>
banded_layout* banded_get(b200_handle* h, edge_partition& ep, const RT* weights, int tile_edges,
                          int row_cap);

// schedule of the reduced partition (graph.cu): rows [0, n_long) packed on the host, the rest
// by edge windows on the device
void build_schedule_reduced(b200_handle* h, edge_partition& rp, int64_t n_long, int64_t n_nz, int T,
                            int rowcap);

}  // namespace b200
