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
constexpr int kMinCell     = 3;      // cells with fewer edges stay direct gathers
constexpr int kCellMax     = 2048;   // cells are cut at CSC positions that are multiples of this
constexpr int kTaskWindow  = 2048;   // a task = the cells that start inside one window of a band
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

// tasks [task_first, task_end) of one band, run by one CTA with the band in shared memory
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
  dbuf<band_group> groups;
  dbuf<int32_t> cta_group_first;  // [num_ctas + 1]
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
  const int32_t* cta_group_first;
  int64_t num_minors;
  const iter_state* state;        // may be nullptr
};

// ------------------------------------------------------------------------------------------
// phase-1 kernel: one persistent CTA per SM.  For each of its groups the CTA loads the band of
// x into shared memory; its 32 warps then claim the group's tasks.  A warp walks a task in
// steps of 256 edges (lane l owns edges [8l, 8l + 8) of the step): one 128-bit load of eight
// 16-bit indices (prefetched a step ahead), eight shared-memory gathers, a sequential
// segmented sum inside the lane, a segmented warp scan of the lanes' open tails, and one store
// per cell that ends in the lane.  The sum that is still open at the end of a step is carried
// in a register to the next step.
// ------------------------------------------------------------------------------------------
template <typename RT, bool WEIGHTED>
__global__ void __launch_bounds__(kBandThreads, 1) band_reduce_kernel(const __grid_constant__ band_params P)
{
  using G = band_geom<RT>;
  extern __shared__ __align__(16) unsigned char band_smem[];
  RT* band = reinterpret_cast<RT*>(band_smem);
  __shared__ int s_next;

  if (P.state && P.state->done) return;
  const int tid = threadIdx.x, lane = tid & 31;
  const RT* x   = static_cast<const RT*>(P.x);
  RT* pval      = static_cast<RT*>(P.pval);
  const RT* bw  = static_cast<const RT*>(P.bw);

  const int g0 = P.cta_group_first[blockIdx.x], g1 = P.cta_group_first[blockIdx.x + 1];
  for (int g = g0; g < g1; ++g) {
    const band_group grp = P.groups[g];
    __syncthreads();  // every warp is done with the previous band
    {
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
      if (tid == 0) s_next = grp.task_first;
    }
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
      int64_t rank       = t.rank_first;   // rank of the next cell start this warp will meet
      RT carry           = RT(0);          // sum of the cell that is open at the step boundary
      uint4 cur          = __ldcs(reinterpret_cast<const uint4*>(ip));
      for (int st = 0; st < steps; ++st) {
        uint4 nxt = make_uint4(0, 0, 0, 0);
        if (st + 1 < steps) nxt = __ldcs(reinterpret_cast<const uint4*>(ip + (st + 1) * kStepEdges));
        const int base = st * kStepEdges + lane * 8;   // position of the lane's first edge
        uint32_t h[8];
        h[0] = cur.x & 0xffffu; h[1] = cur.x >> 16; h[2] = cur.y & 0xffffu; h[3] = cur.y >> 16;
        h[4] = cur.z & 0xffffu; h[5] = cur.z >> 16; h[6] = cur.w & 0xffffu; h[7] = cur.w >> 16;
        uint32_t fmask = 0, vmask = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const bool valid = (base + i >= lead) && (base + i < total);
          vmask |= (valid ? 1u : 0u) << i;
          fmask |= ((valid && (h[i] & 0x8000u)) ? 1u : 0u) << i;
        }
        // ranks of the lane's cell starts
        const int cnt = __popc(fmask);
        int incl      = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int u = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += u;
        }
        const int excl       = incl - cnt;
        const int step_flags = __shfl_sync(0xffffffffu, incl, 31);
        // slot of the cell that ENDS at each of the lane's cell starts (= the previous cell);
        // loaded now, used after the gathers
        int slot[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          slot[i] = -1;
          if (fmask & (1u << i)) {
            const int64_t r = rank + excl + __popc(fmask & ((1u << i) - 1u)) - 1;
            if (r >= t.rank_first) slot[i] = __ldg(P.ppos + r);
          }
        }
        // gathers from the band
        RT v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (vmask & (1u << i)) ? band[h[i] & G::mask] : RT(0);
        if (WEIGHTED) {
          if (vmask) {
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
        }
        // sequential segmented sum inside the lane: closev[i] = sum of the lane's edges since
        // its previous cell start (or since the lane's first edge), taken at cell start i
        RT closev[8];
        RT run = RT(0);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          closev[i] = run;
          if (fmask & (1u << i)) run = RT(0);
          run += v[i];
        }
        // segmented inclusive scan of the open tails across the lanes
        RT sv      = run;
        bool sflag = cnt > 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const RT pv   = __shfl_up_sync(0xffffffffu, sv, o);
          const int pf  = __shfl_up_sync(0xffffffffu, sflag ? 1 : 0, o);
          if (lane >= o) {
            if (!sflag) sv += pv;
            sflag = sflag || (pf != 0);
          }
        }
        // what is open when the lane's first edge is reached
        RT up_v        = __shfl_up_sync(0xffffffffu, sv, 1);
        const int up_f = __shfl_up_sync(0xffffffffu, sflag ? 1 : 0, 1);
        RT carry_in    = carry;
        if (lane > 0) carry_in = up_f ? up_v : up_v + carry;
        const RT last_v   = __shfl_sync(0xffffffffu, sv, 31);
        const int last_f  = __shfl_sync(0xffffffffu, sflag ? 1 : 0, 31);
        // stores: the first cell start of the lane closes the cell carried into the lane
        bool first = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (fmask & (1u << i)) {
            const RT val = first ? closev[i] + carry_in : closev[i];
            first        = false;
            if (slot[i] >= 0) pval[slot[i]] = val;
          }
        }
        carry = last_f ? last_v : last_v + carry;
        rank += step_flags;
        cur = nxt;
      }
      // the cell that is still open at the end of the task
      if (lane == 0 && rank > t.rank_first) pval[__ldg(P.ppos + rank - 1)] = carry;
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
