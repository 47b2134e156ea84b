// The hot path: pull-direction edge->vertex transform-reduce over a CSC rectangle, hand-written
// for sm_100a.  B200-native replacement of the four kernels of
// cpp/src/prims/per_v_transform_reduce_incoming_outgoing_e.cuh (:84 hypersparse, :194 low,
// :306 mid, :401 high) and of their host orchestration (:489-1038).
//
// Design (DESIGN.md "kernels"; the numbers behind it are in profiles/):
//  * the edge arrays are cut ONCE (at graph build) into tiles of <= T edges that respect the
//    reference's degree segments: rows of the high-degree segment (degree >= 1024,
//    graph_view.hpp:255-263) are split into chunks of T edges that many CTAs reduce
//    cooperatively (block-wide reduction + a deterministic combine in the tail kernel);
//    mid- and low-degree rows are packed into tiles of COMPLETE rows.
//  * one persistent, warp-specialised CTA per SM walks the tiles round-robin:
//      - 1 producer thread streams each tile's indices / weights / offsets into a STAGES-deep
//        shared-memory ring with cp.async.bulk (TMA bulk copy, SASS UBLKCP) + mbarriers; the
//        index stream carries an L2 evict-first policy so it does not displace the gathered
//        vector;
//      - 16 gather warps read 4 indices at a time (128-bit LDS), gather x[idx] and write the
//        products back into the ring.  The first `hub` entries of x - the highest-degree
//        sources, which the degree-descending renumbering puts first - are kept in shared
//        memory (measured: 4.3 random gathers/clk/SM from shared memory vs 1.0 through
//        L1TEX, profiles/microbench_gather.md), everything else goes through ld.global.nc;
//      - 15 reduce warps turn products into row sums from shared memory with G = 1..32 lanes
//        per row (thread-per-vertex for the low-degree segment, sub-warp / warp segmented
//        shuffles for the mid-degree segment; G chosen per tile from its mean degree) and run
//        the caller's fused epilogue.
//    The three roles only meet at mbarriers, so the L1TEX gather pipe never waits for a
//    reduction or an epilogue.
//  * the caller's epilogue (PageRank: + unvarying part, new scaled vector, dangling mass and
//    L1 difference; Katz: + beta, L1 difference; raw: y = init + scale*sum) is fused into the
//    row reduction, so one launch per power iteration does what the reference does in
//    ~10 launches and 2 host synchronisations (pagerank_impl.cuh:209-292).
#pragma once

#include "common.cuh"
#include "epilogues.cuh"
#include "prims.cuh"
#include "ptx.cuh"

namespace b200 {

struct pull_params {
  const tile_t* tiles;
  int64_t num_tiles;
  const void* offsets;     // ET [num_rows + 1]
  const int32_t* indices;  // padded by >= 16 bytes
  const void* weights;     // RT or nullptr
  const void* x;           // gathered vector, RT
  // second piece of the gathered vector (banded.cuh): an index >= x_split reads
  // x_hi[index - x_split] (the partial sums of the pre-reduced cells); INT32_MAX = none
  const void* x_hi;
  int32_t x_split;
  int64_t zero_first;      // rows [zero_first, num_rows) have no edges (epilogue only)
  int64_t num_rows;
  double* chunk_partials;  // [num_chunks]
  double* cta_partials;    // [2 (a,b)][slots]  written at [blockIdx.x]
  int32_t slots;
  int32_t hub_entries;     // entries of x cached in shared memory, all pieces together
  // the cached entries are the heads of up to kMaxHubPieces pieces of x (one on a single GPU;
  // one per row_comm peer on a 2D grid, whose minor range concatenates R degree-sorted ranges)
  int32_t hub_pieces;
  int32_t hub_piece_start[4];  // first x index of piece k
  int32_t hub_piece_len[4];    // cached entries of piece k (sum = hub_entries)
  const iter_state* state; // may be nullptr (raw spmv)
};

#ifdef B200_PROFILE_ROLES
// tuning builds only: clocks one lane of each role spends waiting / working, per CTA
//  [0] kernel  [1] gather wait full  [2] gather zero rows  [3] reduce wait full
//  [4] reduce wait ready  [5] producer wait empty  [6] tiles  [7] gather retire stalls (unused)
__device__ unsigned long long b200_role_clocks[256][8];
#define B200_CLK(var) const long long var = clock64()
#define B200_ACC(acc, t0) acc += clock64() - (t0)
#else
#define B200_CLK(var)
#define B200_ACC(acc, t0)
#endif

constexpr int kMaxHubPieces  = 4;
constexpr int kLongRowDegree = 1024;  // = mid_degree_threshold, graph_view.hpp:255-263

// tuning builds override the warp split (scripts/build_variant.sh); 31 worker warps in total
#ifndef B200_GATHER_WARPS
#define B200_GATHER_WARPS 16
#endif
constexpr int kGatherWarps  = B200_GATHER_WARPS;
constexpr int kReduceWarps  = 31 - kGatherWarps;
constexpr int kPullThreads  = 1024;  // gather + reduce + 1 producer warp
constexpr int kGatherThreads = kGatherWarps * 32;
constexpr int kReduceThreads = kReduceWarps * 32;

// shared-memory geometry of one CTA (dynamic part): [hub | ring stages | product ring]
// ROWCAP = most rows the schedule puts into one tile (their offsets and epilogue inputs are staged)
template <typename ET, typename RT, bool WEIGHTED, int T, int STAGES, int ROWCAP, int NSTAGED>
struct pull_smem {
  // staged element capacity: T edges + one row overhang (< 1024) + alignment slack
  static constexpr int cap       = T + kLongRowDegree + 8;
  static constexpr size_t idx_b  = size_t(cap) * 4;
  static constexpr size_t w_b    = WEIGHTED ? size_t(cap) * sizeof(RT) : 0;
  static constexpr size_t off_b  = size_t(ROWCAP + 8) * sizeof(ET);
  static constexpr size_t in_b   = size_t(ROWCAP + 8) * sizeof(RT);  // one staged epilogue input
  static constexpr size_t desc_b = 32;  // the tile descriptor travels with the tile (TMA)
  static constexpr size_t stage  = ((idx_b + w_b + off_b + NSTAGED * in_b + desc_b) + 127) & ~size_t(127);
  // 4-byte results overwrite the staged indices in place; 8-byte results get their own ring
  static constexpr bool inplace  = sizeof(RT) == 4;
  static constexpr size_t prod_b = inplace ? 0 : ((size_t(cap) * sizeof(RT) + 127) & ~size_t(127));
  static constexpr size_t ring   = STAGES * (stage + prod_b);
  static constexpr size_t max_dyn = 232448 - 2048;  // 227 KB minus the static part
  static constexpr int max_hub   = static_cast<int>((max_dyn - ring) / sizeof(RT)) & ~31;
  static size_t total(int hub) { return ring + ((size_t(hub) * sizeof(RT) + 127) & ~size_t(127)); }
};

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <typename ET, typename RT, bool WEIGHTED, typename EPI, int T, int STAGES, int ROWCAP>
__global__ void __launch_bounds__(kPullThreads, 1)
  pull_tiles_kernel(const __grid_constant__ pull_params P, EPI epi)
{
  using S = pull_smem<ET, RT, WEIGHTED, T, STAGES, ROWCAP, EPI::kStage>;
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t full_bar[STAGES];   // TMA landed            (1 arrival + tx bytes)
  __shared__ uint64_t ready_bar[STAGES];  // products written      (kGatherWarps arrivals)
  __shared__ uint64_t empty_bar[STAGES];  // stage consumed        (kReduceWarps arrivals)
  __shared__ double warp_part[STAGES][kGatherWarps];
  __shared__ double red_a[32], red_b[32];

  if (P.state && P.state->done) return;
#ifdef B200_PROFILE_ROLES
  const long long prof_t0 = clock64();
  long long prof_a = 0, prof_b = 0;
#endif

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int hub_n           = P.hub_entries;
  RT* hub                   = reinterpret_cast<RT*>(smem);
  unsigned char* ring_base  = smem + ((size_t(hub_n) * sizeof(RT) + 127) & ~size_t(127));
  unsigned char* prod_base  = ring_base + STAGES * S::stage;

  auto idx_s = [&](int s) { return reinterpret_cast<int32_t*>(ring_base + s * S::stage); };
  auto w_s   = [&](int s) { return reinterpret_cast<RT*>(ring_base + s * S::stage + S::idx_b); };
  auto off_s = [&](int s) {
    return reinterpret_cast<ET*>(ring_base + s * S::stage + S::idx_b + S::w_b);
  };
  auto in_s = [&](int s, int k) {
    return reinterpret_cast<RT*>(ring_base + s * S::stage + S::idx_b + S::w_b + S::off_b +
                                 k * S::in_b);
  };
  auto desc_s = [&](int s) {
    return reinterpret_cast<tile_t*>(ring_base + s * S::stage + S::idx_b + S::w_b + S::off_b +
                                     EPI::kStage * S::in_b);
  };
  auto prod_s = [&](int s) {
    return S::inplace ? reinterpret_cast<RT*>(ring_base + s * S::stage)
                      : reinterpret_cast<RT*>(prod_base + s * S::prod_b);
  };

  const ET* offsets    = static_cast<const ET*>(P.offsets);
  const RT* weights    = static_cast<const RT*>(P.weights);
  const RT* x          = static_cast<const RT*>(P.x);
  const int x_split    = P.x_split;
  // pre-offset so that both pieces are addressed with the index itself
  const RT* x_hi       = static_cast<const RT*>(P.x_hi) - (P.x_hi ? x_split : 0);
  const int64_t stride = gridDim.x;
  const int64_t my_n =
    (P.num_tiles > blockIdx.x) ? (P.num_tiles - blockIdx.x + stride - 1) / stride : 0;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&ready_bar[s], kGatherWarps);
      mbar_init(&empty_bar[s], kReduceWarps);
    }
    fence_mbar_init();
  }
  // hub cache: the highest-degree sources live at the front of every piece of x
  // (degree-descending ids)
  {
    int base = 0;
    for (int k = 0; k < P.hub_pieces; ++k) {
      const RT* xp = x + P.hub_piece_start[k];
      for (int i = tid; i < P.hub_piece_len[k]; i += kPullThreads) hub[base + i] = xp[i];
      base += P.hub_piece_len[k];
    }
  }
  epi.setup(P.state);
  __syncthreads();

  acc2 acc;

  if (warp == kGatherWarps + kReduceWarps) {
    // ================= producer: one thread drives the TMA ring =================
    if (lane == 0) {
      const uint64_t pol_stream = policy_evict_first();
      const uint64_t pol_keep   = policy_evict_last();
      // Per tile: what to copy and from where.  The descriptor of tile j + 1 is read while tile
      // j's copies are issued, and every descriptor is also written into its stage, so the
      // other warps never touch global memory for it.
      struct span_t {
        int64_t a0, r0, q0;
        uint32_t cnt, off_bytes, in_bytes;
      };
      auto spans = [&](tile_t const& t) {
        span_t sp{};
        const int lead = static_cast<int>(t.edge_first & 3);
        sp.a0          = t.edge_first - lead;
        sp.cnt         = static_cast<uint32_t>((lead + t.num_edges + 3) & ~3);
        if (t.kind == 0) {
          constexpr int per16 = 16 / sizeof(ET);
          sp.r0               = t.row_first & ~int64_t(per16 - 1);
          int n_off           = static_cast<int>(t.row_first - sp.r0) + t.num_rows + 1;
          n_off               = (n_off + per16 - 1) & ~(per16 - 1);
          sp.off_bytes        = static_cast<uint32_t>(n_off * sizeof(ET));
          if (EPI::kStage > 0) {
            constexpr int q16 = 16 / sizeof(RT);
            sp.q0             = t.row_first & ~int64_t(q16 - 1);
            int n_in          = static_cast<int>(t.row_first - sp.q0) + t.num_rows;
            n_in              = (n_in + q16 - 1) & ~(q16 - 1);
            sp.in_bytes       = static_cast<uint32_t>(n_in * sizeof(RT));
          }
        }
        return sp;
      };
      // (an L2 bulk prefetch several tiles ahead was tried and measured neutral to negative:
      // the ring is not bound by DRAM latency but by per-operation issue cost, so the
      // producer issues as few asynchronous operations per tile as possible)
      auto tile_at = [&](int64_t j) { return P.tiles[blockIdx.x + j * stride]; };
      tile_t t = my_n > 0 ? tile_at(0) : tile_t{};
      for (int64_t j = 0; j < my_n; ++j) {
        const int s = static_cast<int>(j % STAGES);
        tile_t tn = t;
        if (j + 1 < my_n) tn = tile_at(j + 1);
        {
          B200_CLK(w0);
          if (j >= STAGES) mbar_wait(&empty_bar[s], static_cast<uint32_t>(((j / STAGES) - 1) & 1));
          B200_ACC(prof_a, w0);
        }
        fence_proxy_async();
        const span_t sp = spans(t);
        uint32_t bytes  = sp.cnt * 4u;
        if (WEIGHTED) bytes += sp.cnt * static_cast<uint32_t>(sizeof(RT));
        // the descriptor goes into the stage with plain stores; the arrive below releases them
        // to every thread that acquires the barrier phase
        *desc_s(s) = t;
        mbar_arrive_expect_tx(&full_bar[s], bytes + sp.off_bytes + EPI::kStage * sp.in_bytes);
        if (sp.cnt) bulk_g2s(idx_s(s), P.indices + sp.a0, sp.cnt * 4u, &full_bar[s], pol_stream);
        if (WEIGHTED && sp.cnt)
          bulk_g2s(w_s(s), weights + sp.a0, sp.cnt * static_cast<uint32_t>(sizeof(RT)), &full_bar[s],
                   pol_stream);
        if (sp.off_bytes) bulk_g2s(off_s(s), offsets + sp.r0, sp.off_bytes, &full_bar[s], pol_stream);
        if (sp.in_bytes) {
#pragma unroll
          for (int k = 0; k < EPI::kStage; ++k)
            bulk_g2s(in_s(s, k), epi.stage_src(k) + sp.q0, sp.in_bytes, &full_bar[s], pol_keep);
        }
        t = tn;
      }
#ifdef B200_PROFILE_ROLES
      b200_role_clocks[blockIdx.x][5] = prof_a;
      b200_role_clocks[blockIdx.x][6] = my_n;
#endif
    }
  } else if (warp < kGatherWarps) {
    // ================= gather warps =================
    const int hub0 = P.hub_piece_len[0];
    const bool multi_piece = P.hub_pieces > 1;
    auto hub_slot = [&](int id) -> int {
      // slot of x[id] in the shared-memory cache, or -1
      if (id < hub0) return id;
      if (multi_piece) {
        int base = hub0;
#pragma unroll
        for (int k = 1; k < kMaxHubPieces; ++k) {
          if (k < P.hub_pieces) {
            const unsigned d = static_cast<unsigned>(id - P.hub_piece_start[k]);
            if (d < static_cast<unsigned>(P.hub_piece_len[k])) return base + static_cast<int>(d);
            base += P.hub_piece_len[k];
          }
        }
      }
      return -1;
    };
#ifdef B200_EXP_NOGATHER  // experiment: what the kernel costs without the global gathers
    auto gather1 = [&](int id) -> RT { const int sl = hub_slot(id); return sl >= 0 ? hub[sl] : RT(1); };
#else
    // Tried and measured slower (profiles/r1_sweep_l1_hot.log, r1_notes.md): no-allocate loads
    // for the cold tail of x (-15 %); the zero-row epilogues interleaved into the reduce warps
    // instead of after the gather warps' tiles (-9 %: the reduce path is on the ring's
    // critical path, the gather warps' tail is not).
    auto gather1 = [&](int id) -> RT {
      const int sl = hub_slot(id);
      return sl >= 0 ? hub[sl] : ld_gather((id >= x_split ? x_hi : x) + id);
    };
#endif
    // Rolling software pipeline: a lane issues the four gathers of its next quad (possibly of
    // the NEXT tile) before it retires the previous quad, so the L1TEX queue never drains at a
    // tile boundary.  A pending quad carries the stage it belongs to; the warp arrives on that
    // stage's `ready` barrier when it retires the tile's last quad.
    constexpr int kStep = kGatherThreads * 4;
    constexpr int kIter = (S::cap + kStep - 1) / kStep;
    constexpr unsigned kActive = 1u, kLast = 2u, kChunk = 4u, kExists = 8u;
    RT pv0 = RT(0), pv1 = RT(0), pv2 = RT(0), pv3 = RT(0);
    int p_i = 0, p_s = 0;
    unsigned p_flags = 0;
    double part = 0.0;
    auto retire = [&]() {
      if (!(p_flags & kExists)) return;
      const int s = p_s;
      if (p_flags & kActive) {
        RT v0 = pv0, v1 = pv1, v2 = pv2, v3 = pv3;
        if (WEIGHTED) {
          const RT* ws = w_s(s);
          v0 *= ws[p_i]; v1 *= ws[p_i + 1]; v2 *= ws[p_i + 2]; v3 *= ws[p_i + 3];
        }
        if (p_flags & kChunk) {
          // chunk of a long row: block-wide reduction, no shared-memory round trip
          const tile_t tp = *desc_s(s);
          const int lead  = static_cast<int>(tp.edge_first & 3);
          const int hi    = lead + tp.num_edges;
          RT sum4 = RT(0);
          if (p_i >= lead && p_i < hi) sum4 += v0;
          if (p_i + 1 >= lead && p_i + 1 < hi) sum4 += v1;
          if (p_i + 2 >= lead && p_i + 2 < hi) sum4 += v2;
          if (p_i + 3 >= lead && p_i + 3 < hi) sum4 += v3;
          part += static_cast<double>(sum4);
        } else {
          RT* ps = prod_s(s) + p_i;
          if constexpr (sizeof(RT) == 4) {
            *reinterpret_cast<float4*>(ps) = make_float4(v0, v1, v2, v3);
          } else {
            *reinterpret_cast<double2*>(ps)     = make_double2(v0, v1);
            *reinterpret_cast<double2*>(ps + 2) = make_double2(v2, v3);
          }
        }
      }
      if (p_flags & kLast) {
        if (p_flags & kChunk) {
          double tot = warp_sum(part);
          if (lane == 0) warp_part[s][warp] = tot;
          part = 0.0;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&ready_bar[s]);
      }
      p_flags = 0;
    };
    for (int64_t j = 0; j < my_n; ++j) {
      const int s       = static_cast<int>(j % STAGES);
      const uint32_t ph = static_cast<uint32_t>((j / STAGES) & 1);
      {
        B200_CLK(w0);
        mbar_wait(&full_bar[s], ph);
        B200_ACC(prof_a, w0);
      }
      const tile_t t    = *desc_s(s);
      const int lead    = static_cast<int>(t.edge_first & 3);
      const int cnt     = (lead + t.num_edges + 3) & ~3;
      const int32_t* is = idx_s(s);
      const unsigned tile_flags = kExists | (t.kind != 0 ? kChunk : 0u);
#pragma unroll
      for (int u = 0; u < kIter; ++u) {
        const int i    = tid * 4 + u * kStep;
        const bool act = i < cnt;
        RT n0 = RT(0), n1 = RT(0), n2 = RT(0), n3 = RT(0);
        if (act) {
          const int4 id = *reinterpret_cast<const int4*>(is + i);
          n0 = gather1(id.x); n1 = gather1(id.y); n2 = gather1(id.z); n3 = gather1(id.w);
        }
        retire();
        pv0 = n0; pv1 = n1; pv2 = n2; pv3 = n3;
        p_i = i; p_s = s;
        p_flags = tile_flags | (act ? kActive : 0u) | (u == kIter - 1 ? kLast : 0u);
      }
    }
    retire();
    // rows without edges: epilogue only (per_v...cuh:553-559 fills them with init)
    B200_CLK(z0);
    {
      const int64_t step = static_cast<int64_t>(gridDim.x) * kGatherThreads;
      int64_t r = P.zero_first + static_cast<int64_t>(blockIdx.x) * kGatherThreads + tid;
      for (; r + 3 * step < P.num_rows; r += 4 * step) {
        typename EPI::pre_t p0 = epi.prefetch(r), p1 = epi.prefetch(r + step),
                            p2 = epi.prefetch(r + 2 * step), p3 = epi.prefetch(r + 3 * step);
        epi.finish(r, RT(0), p0, acc);
        epi.finish(r + step, RT(0), p1, acc);
        epi.finish(r + 2 * step, RT(0), p2, acc);
        epi.finish(r + 3 * step, RT(0), p3, acc);
      }
      for (; r < P.num_rows; r += step) epi(r, RT(0), acc);
    }
#ifdef B200_PROFILE_ROLES
    B200_ACC(prof_b, z0);
    if (tid == 0) {
      b200_role_clocks[blockIdx.x][1] = prof_a;
      b200_role_clocks[blockIdx.x][2] = prof_b;
    }
#endif
  } else {
    // ================= reduce warps =================
    const int rtid  = tid - kGatherThreads;
    const int rwarp = warp - kGatherWarps;
    for (int64_t j = 0; j < my_n; ++j) {
      const int s       = static_cast<int>(j % STAGES);
      const uint32_t ph = static_cast<uint32_t>((j / STAGES) & 1);
      {
        B200_CLK(w0);
        mbar_wait(&full_bar[s], ph);   // descriptor + offsets landed (TMA)
        B200_ACC(prof_a, w0);
      }
      const tile_t t    = *desc_s(s);
      const int lead    = static_cast<int>(t.edge_first & 3);
      {
        B200_CLK(w1);
        mbar_wait(&ready_bar[s], ph);  // products written (gather warps)
        B200_ACC(prof_b, w1);
      }
#ifdef B200_EXP_NOREDUCE  // experiment: what the kernel costs without the row reductions
      if (false) {
#else
      if (t.kind == 0) {
#endif
        RT* ps            = prod_s(s);
        constexpr int per16 = 16 / sizeof(ET);
        constexpr int q16   = 16 / sizeof(RT);
        const ET* offp      = off_s(s) + (t.row_first & (per16 - 1));
        const RT* in0       = in_s(s, 0) + (t.row_first & (q16 - 1));
        const RT* in1       = in_s(s, EPI::kStage > 1 ? 1 : 0) + (t.row_first & (q16 - 1));
        const int64_t a0 = t.edge_first - lead;
        const int nrows  = t.num_rows;
        const int avg    = t.num_edges / (nrows > 0 ? nrows : 1);
        // Step 1: row sums with G lanes per row, kU rows of a lane group in flight at once (the
        // reduction is a chain of dependent shared-memory loads and shuffles: without the
        // interleave it was latency-bound at ~2500 clocks per tile).  The sum of a non-empty
        // row is left in the row's own first product slot.
        auto reduce_rows = [&](auto G_) {
          constexpr int G      = decltype(G_)::value;
          constexpr int groups = kReduceThreads / G;
          constexpr int kU     = 4;
          const int gid = rtid / G, l = rtid % G;
          const int rounds = (nrows + groups - 1) / groups;
          for (int k0 = 0; k0 < rounds; k0 += kU) {
            int b[kU], e[kU], i[kU];
            RT sum[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) {
              const int r = (k0 + u) * groups + gid;
              b[u] = 0; e[u] = 0;
              if (r < nrows) {
                b[u] = static_cast<int>(static_cast<int64_t>(offp[r]) - a0);
                e[u] = static_cast<int>(static_cast<int64_t>(offp[r + 1]) - a0);
              }
              i[u]   = b[u] + l;
              sum[u] = RT(0);
            }
            bool more = true;
            while (more) {
              more = false;
#pragma unroll
              for (int u = 0; u < kU; ++u) {
                if (i[u] < e[u]) {
                  sum[u] += ps[i[u]];
                  i[u] += G;
                  more = more || (i[u] < e[u]);
                }
              }
            }
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) {
#pragma unroll
              for (int u = 0; u < kU; ++u) sum[u] += __shfl_xor_sync(0xffffffffu, sum[u], o, G);
            }
            if (l == 0) {
#pragma unroll
              for (int u = 0; u < kU; ++u)
                if (b[u] < e[u]) ps[b[u]] = sum[u];
            }
          }
        };
        if (avg < 3) reduce_rows(std::integral_constant<int, 1>{});
        else if (avg < 6) reduce_rows(std::integral_constant<int, 2>{});
        else if (avg < 12) reduce_rows(std::integral_constant<int, 4>{});
        else if (avg < 24) reduce_rows(std::integral_constant<int, 8>{});
        else if (avg < 48) reduce_rows(std::integral_constant<int, 16>{});
        else reduce_rows(std::integral_constant<int, 32>{});
        // Step 2: the epilogue, one thread per row: independent iterations, coalesced stores
        named_bar_sync(1, kReduceThreads);
        for (int r = rtid; r < nrows; r += kReduceThreads) {
          const int b = static_cast<int>(static_cast<int64_t>(offp[r]) - a0);
          const int e = static_cast<int>(static_cast<int64_t>(offp[r + 1]) - a0);
          const RT sum = (b < e) ? ps[b] : RT(0);
          const int64_t row = static_cast<int64_t>(t.row_first) + r;
          epi.finish(row, sum, epi.from_staged(in0, in1, r, row), acc);
        }
      } else if (rwarp == 0) {
        double v = (lane < kGatherWarps) ? warp_part[s][lane] : 0.0;
        v        = warp_sum(v);
        if (lane == 0) P.chunk_partials[t.aux] = v;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
#ifdef B200_PROFILE_ROLES
    if (tid == kGatherThreads) {
      b200_role_clocks[blockIdx.x][3] = prof_a;
      b200_role_clocks[blockIdx.x][4] = prof_b;
    }
#endif
  }

#ifdef B200_PROFILE_ROLES
  __syncthreads();
  if (tid == 0) b200_role_clocks[blockIdx.x][0] = clock64() - prof_t0;
#endif
  // per-CTA partial sums of the epilogue scalars (fixed order: deterministic)
  if (P.cta_partials) {
    double a = warp_sum(acc.a), b = warp_sum(acc.b);
    if (lane == 0) { red_a[warp] = a; red_b[warp] = b; }
    __syncthreads();
    if (warp == 0) {
      a = warp_sum(red_a[lane]);
      b = warp_sum(red_b[lane]);
      if (lane == 0) {
        P.cta_partials[blockIdx.x]           = a;
        P.cta_partials[P.slots + blockIdx.x] = b;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// tail kernel: combine the chunk partials of the long rows (deterministic order), run their
// epilogue, then the last block to finish folds all per-CTA partial sums into the iteration
// state and applies the driver's stop test.
// ------------------------------------------------------------------------------------------
struct tail_params {
  const int32_t* long_row_ids;          // [n_long] row of every long row, or nullptr = identity
  const int64_t* long_row_chunk_first;  // [n_long + 1]
  const double* chunk_partials;
  int64_t num_long_rows;
  double* cta_partials;   // [2][slots]
  int32_t slots;
  int32_t main_ctas;      // slots [0, main_ctas) belong to the main kernel
  iter_state* state;      // may be nullptr (raw)
  double epsilon;         // stop when b < epsilon (b * 1 for pagerank/katz)
  int64_t iteration;      // 1-based index of the iteration this launch completes
  int32_t check_convergence;
  int32_t finalize;       // fold partials into state
};

template <typename RT, typename EPI, int THREADS>
__global__ void __launch_bounds__(THREADS) pull_tail_kernel(const __grid_constant__ tail_params P,
                                                            EPI epi)
{
  __shared__ double red_scratch[32];
  __shared__ bool is_last;
  if (P.state && P.state->done) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  epi.setup(P.state);
  acc2 acc;
  // 8 lanes per long row (most long rows have a handful of chunks); fixed order => deterministic
  constexpr int G           = 8;
  const int64_t groups_total = static_cast<int64_t>(gridDim.x) * (THREADS / G);
  const int l               = tid % G;
  const int64_t rounds      = (P.num_long_rows + groups_total - 1) / groups_total;
  for (int64_t k = 0; k < rounds; ++k) {
    const int64_t r = k * groups_total + static_cast<int64_t>(blockIdx.x) * (THREADS / G) + tid / G;
    const bool act  = r < P.num_long_rows;
    double s        = 0.0;
    int64_t row     = r;
    typename EPI::pre_t pre{};
    if (act) {
      const int64_t c0 = P.long_row_chunk_first[r], c1 = P.long_row_chunk_first[r + 1];
      if (P.long_row_ids) row = P.long_row_ids[r];
      if (l == 0) pre = epi.prefetch(row);
      for (int64_t c = c0 + l; c < c1; c += G) s += P.chunk_partials[c];
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o, G);
    if (act && l == 0) epi.finish(row, static_cast<RT>(s), pre, acc);
  }
  (void)lane; (void)warp;
  if (!P.cta_partials) return;
  double a = block_sum(acc.a, red_scratch);
  double b = block_sum(acc.b, red_scratch);
  if (tid == 0) {
    P.cta_partials[P.main_ctas + blockIdx.x]           = a;
    P.cta_partials[P.slots + P.main_ctas + blockIdx.x] = b;
  }
  if (!P.finalize) return;
  __threadfence();
  if (tid == 0) {
    uint32_t t = atomicAdd(&P.state->ticket, 1u);
    is_last    = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  const int n = P.main_ctas + static_cast<int>(gridDim.x);
  double sa = 0.0, sb = 0.0;
  const volatile double* cp = P.cta_partials;
  for (int i = tid; i < n; i += THREADS) {
    sa += cp[i];
    sb += cp[P.slots + i];
  }
  sa = block_sum(sa, red_scratch);
  sb = block_sum(sb, red_scratch);
  if (tid == 0) {
    iter_state* st = P.state;
    st->a          = sa;
    st->b          = sb;
    st->iterations = P.iteration;
    st->ticket     = 0;
    if (P.check_convergence && sb < P.epsilon) {
      st->converged = 1;
      st->done      = 1;
    }
  }
}

}  // namespace b200
