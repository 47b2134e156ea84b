// Power-iteration drivers on top of the pull kernel (spmv.cuh):
//   PageRank     cpp/src/link_analysis/pagerank_impl.cuh:48-293
//   Katz         cpp/src/centrality/katz_centrality_impl.cuh:42-173
//   eigenvector  cpp/src/centrality/eigenvector_centrality_impl.cuh:45-138
// One SpMV launch + one tail launch per iteration; dangling mass, L1 difference and the stop
// test live on the device (iter_state), the host only polls the state every few iterations
// instead of synchronising twice per iteration as the reference does (pagerank_impl.cuh:218,278).
#include <cmath>
#include <cstdlib>

#include <climits>

#include <array>
#include <atomic>

#include "banded.cuh"
#include "sell.cuh"
#include "centrality.cuh"
#include "spmv.cuh"

namespace b200 {

namespace {

constexpr int kVThreads = 256;

inline unsigned vgrid(size_t n, int sm_count)
{
  size_t b = (n + kVThreads * 4 - 1) / (kVThreads * 4);
  size_t cap = static_cast<size_t>(sm_count) * 8;
  if (b > cap) b = cap;
  if (b == 0) b = 1;
  return static_cast<unsigned>(b);
}

// ---- kernel configurations -----------------------------------------------------------------
// T = edges per tile, STAGES = depth of the TMA ring, ROWCAP = most rows per tile
template <int T_, int STAGES_, int ROWCAP_>
struct pull_cfg {
  static constexpr int T      = T_;
  static constexpr int STAGES = STAGES_;
  static constexpr int ROWCAP = ROWCAP_;
};
template <typename RT, bool WEIGHTED>
struct default_cfg {
  using type = pull_cfg<4096, 4, 512>;  // f32 unweighted (best of the sweep in profiles/)
};
template <>
struct default_cfg<float, true> {
  using type = pull_cfg<2048, 4, 512>;
};
template <>
struct default_cfg<double, false> {
  using type = pull_cfg<2048, 4, 512>;
};
template <>
struct default_cfg<double, true> {
  using type = pull_cfg<1024, 4, 256>;
};
// phase 2 of the banded path walks the reduced partition: a quarter of the entries on the same
// rows, so tiles are row-limited and a larger row cap pays (profiles/r1e_sweep_*.log)
template <typename RT, bool WEIGHTED>
struct banded_cfg {
  using type = pull_cfg<4096, 4, 1024>;  // f32 unweighted
};
template <>
struct banded_cfg<float, true> {
  using type = pull_cfg<2048, 4, 1024>;
};
template <>
struct banded_cfg<double, false> {
  using type = pull_cfg<2048, 4, 512>;   // a larger row cap measured slower (r1g_quick_*.log)
};
template <>
struct banded_cfg<double, true> {
  using type = pull_cfg<1024, 4, 256>;
};
// alternatives for tuning sweeps (B200_PULL_CFG=1..), f32 unweighted int32 offsets only
template <int ID>
struct alt_cfg {
  using type = pull_cfg<4096, 4, 512>;
};
template <> struct alt_cfg<1> { using type = pull_cfg<4096, 3, 1024>; };
template <> struct alt_cfg<2> { using type = pull_cfg<4096, 4, 1024>; };
template <> struct alt_cfg<3> { using type = pull_cfg<4096, 3, 512>; };
template <> struct alt_cfg<4> { using type = pull_cfg<4096, 5, 512>; };
template <> struct alt_cfg<5> { using type = pull_cfg<8192, 2, 1024>; };
template <> struct alt_cfg<6> { using type = pull_cfg<4096, 3, 2048>; };
template <> struct alt_cfg<7> { using type = pull_cfg<4096, 4, 2048>; };
template <> struct alt_cfg<8> { using type = pull_cfg<8192, 3, 2048>; };
template <> struct alt_cfg<9> { using type = pull_cfg<2048, 6, 1024>; };
template <> struct alt_cfg<10> { using type = pull_cfg<4096, 5, 1024>; };
constexpr int kNumAltCfg = 11;

// the two-phase banded path (banded.cuh) is the single-GPU default (measured 1.8x the tiled
// kernel alone at scale 24, profiles/r1d_*); B200_BANDED=0 selects the single-phase kernel
constexpr int kBandedDefault = 1;

int env_int(const char* name, int dflt)
{
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : dflt;
}

// fallback for graphs without degree segments (renumber = false): one warp per row straight
// from global memory (what the reference does with its low-degree kernel over all vertices,
// per_v...cuh:814-828, but warp- instead of thread-per-row)
template <typename ET, typename RT, bool WEIGHTED, typename EPI>
__global__ void __launch_bounds__(256) pull_rows_generic_kernel(const ET* offsets,
                                                                const int32_t* indices,
                                                                const RT* weights, const RT* x,
                                                                int64_t num_rows, EPI epi,
                                                                const iter_state* state,
                                                                double* cta_partials, int slots)
{
  __shared__ double red_scratch[32];
  if (state && state->done) return;
  epi.setup(state);
  acc2 acc;
  const int lane      = threadIdx.x & 31;
  const int64_t warp  = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int64_t nwarp = (int64_t(gridDim.x) * blockDim.x) >> 5;
  for (int64_t r = warp; r < num_rows; r += nwarp) {
    int64_t b = offsets[r], e = offsets[r + 1];
    RT sum = RT(0);
    for (int64_t i = b + lane; i < e; i += 32) {
      RT v = __ldg(x + indices[i]);
      if (WEIGHTED) v *= weights[i];
      sum += v;
    }
    sum = warp_sum(sum);
    if (lane == 0) epi(r, sum, acc);
  }
  if (cta_partials) {
    double a = block_sum(acc.a, red_scratch);
    double b = block_sum(acc.b, red_scratch);
    if (threadIdx.x == 0) {
      cta_partials[blockIdx.x]         = a;
      cta_partials[slots + blockIdx.x] = b;
    }
  }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// pull_launcher: everything needed to launch one SpMV (+ tail) over one edge partition
// ------------------------------------------------------------------------------------------
template <typename RT>
struct pull_launcher {
  b200_handle* h;
  edge_partition* ep;
  bool weighted;
  const RT* weights;  // RT-typed (converted copy if the graph stores another type)
  dbuf<char> weights_converted;
  dbuf<double> chunk_partials;
  dbuf<double> cta_partials;
  int main_ctas{0};
  int tail_ctas{1};
  int slots{0};
  int T{0};
  int rowcap{0};
  int cfg_id{0};
  bool tiled{false};
  bool banded_tiles{false};      // phase 2 uses banded_cfg instead of default_cfg
  int64_t launches{0};
  // two-phase path (banded.cuh): phase 1 pre-reduces the cells of the hot bands from shared
  // memory into `pval`, phase 2 runs the tiled kernel over the reduced partition `kp`
  banded_layout* banded{nullptr};
  edge_partition* kp{nullptr};   // the partition the tiled kernel walks (ep or the reduced one)
  dbuf<RT> pval;
  const RT* kweights{nullptr};   // weights of kp
  int band_kernel{2};            // phase-1 formulation: 2 = scratch-column kernel, 1 = first one
  dbuf<int32_t> band_queue;      // two unit counters of phase 1, used alternately
  int64_t band_launches{0};
  // round-2 layout (sell.cuh): both phases as sliced-ELL kernels; replaces `banded` when enabled
  sell_layout* sell{nullptr};
  int sell_ctas1{0};             // phase-1 grid (one CTA per SM)
  int64_t queue_uses{0};         // rotating pool of 8 work-queue counters (sell kernels)
  // Multi-GPU overlap: the minors [split_lo, split_hi) of x are valid when spmv() is called, the
  // rest once `split_event` has fired (the peer copies of the gathered vector run on a second
  // stream).  Phase 1 then runs the bands inside that range first.  Cleared by every spmv().
  int64_t split_lo{0}, split_hi{0};
  cudaEvent_t split_event{nullptr};

  void init(b200_handle* h_, b200_graph* g, edge_partition* ep_);

  template <typename EPI>
  void spmv(const RT* x, EPI const& epi, const iter_state* state, bool want_partials);
  template <typename EPI>
  void tail(EPI const& epi, iter_state* state, double epsilon, int64_t iteration, bool check,
            bool finalize);
};

namespace {
template <typename A, typename B>
__global__ void convert_kernel(const A* in, B* out, size_t n)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<B>(in[i]);
}
}  // namespace

template <typename RT>
void pull_launcher<RT>::init(b200_handle* h_, b200_graph* g, edge_partition* ep_)
{
  h         = h_;
  ep        = ep_;
  weighted  = g->weighted;
  weights   = nullptr;
  cudaStream_t s = h->stream;
  if (weighted) {
    int want = sizeof(RT) == 4 ? B200_FLOAT32 : B200_FLOAT64;
    if (g->weight_type == want) {
      weights = reinterpret_cast<const RT*>(ep->weights.data());
    } else {
      size_t ne = static_cast<size_t>(ep->num_edges);
      weights_converted.alloc(ne * sizeof(RT), s, 32);
      if (ne) {
        if (g->weight_type == B200_FLOAT32)
          convert_kernel<float, RT><<<vgrid(ne, h->sm_count), kVThreads, 0, s>>>(
            reinterpret_cast<const float*>(ep->weights.data()),
            reinterpret_cast<RT*>(weights_converted.data()), ne);
        else
          convert_kernel<double, RT><<<vgrid(ne, h->sm_count), kVThreads, 0, s>>>(
            reinterpret_cast<const double*>(ep->weights.data()),
            reinterpret_cast<RT*>(weights_converted.data()), ne);
        CUDA_TRY(cudaGetLastError());
      }
      weights = reinterpret_cast<const RT*>(weights_converted.data());
    }
  }
  tiled = ep->has_segments;
  cfg_id = 0;
  if (weighted) {
    T      = default_cfg<RT, true>::type::T;
    rowcap = default_cfg<RT, true>::type::ROWCAP;
  } else {
    T      = default_cfg<RT, false>::type::T;
    rowcap = default_cfg<RT, false>::type::ROWCAP;
  }
  if (sizeof(RT) == 4 && !weighted && !ep->offsets64) {
    cfg_id = std::max(0, std::min(kNumAltCfg - 1, env_int("B200_PULL_CFG", 0)));
    const int ts[kNumAltCfg] = {alt_cfg<0>::type::T, alt_cfg<1>::type::T, alt_cfg<2>::type::T,
                                alt_cfg<3>::type::T, alt_cfg<4>::type::T, alt_cfg<5>::type::T,
                                alt_cfg<6>::type::T, alt_cfg<7>::type::T, alt_cfg<8>::type::T,
                                alt_cfg<9>::type::T, alt_cfg<10>::type::T};
    const int rs[kNumAltCfg] = {alt_cfg<0>::type::ROWCAP, alt_cfg<1>::type::ROWCAP,
                                alt_cfg<2>::type::ROWCAP, alt_cfg<3>::type::ROWCAP,
                                alt_cfg<4>::type::ROWCAP, alt_cfg<5>::type::ROWCAP,
                                alt_cfg<6>::type::ROWCAP, alt_cfg<7>::type::ROWCAP,
                                alt_cfg<8>::type::ROWCAP, alt_cfg<9>::type::ROWCAP,
                                alt_cfg<10>::type::ROWCAP};
    T      = ts[cfg_id];
    rowcap = rs[cfg_id];
  }
  kp       = ep;
  kweights = weights;
  banded   = nullptr;
  banded_tiles = false;
  // multi-GPU: every edge partition gets its own layout (bands of its minor range); the
  // opt-in fused rectangle keeps the single-phase kernel
  const bool is_fused = !g->fused.empty() && ep == &g->fused[0];
  sell = nullptr;
  // the sell layout needs no degree order (its rows are window-sorted by reduced length), so
  // graphs built with renumber = false (the HITS companions) take it too
  // Which layout: round 2's sell layout (2) everywhere except on single-GPU graphs below 2^25
  // edges, where round 1's banded path (1) is faster (scale 20 f32: 238 vs 199 GTEPS, f64 181 vs
  // 169; from scale 21 on sell wins: 320 vs 306, scale 22 f64 333 vs 238; profiles/r2_notes.md).
  // B200_LAYOUT overrides.
  int layout = env_int("B200_LAYOUT", -1);
  if (layout < 0)
    layout = (!g->multi_gpu && tiled && ep->num_edges < (int64_t{1} << 25)) ? 1 : 2;
  if (!ep->offsets64 && env_int("B200_BANDED", kBandedDefault) != 0 && layout == 2) {
    // partitions of a multi-GPU graph are only ever run with the raw epilogue into a buffer the
    // caller clears (partition_spmv_run): their all-empty row slices are not stored
    sell_layout* sl = sell_get<RT>(h, *ep, weights, g->multi_gpu);
    if (sl && sl->enabled) {
      sell = sl;
      pval.alloc(static_cast<size_t>(sl->num_partials) + 1, s, 8);
      // the entry behind the last partial is the zero the padding slots of phase 2 read
      CUDA_TRY(cudaMemsetAsync(pval.data() + sl->num_partials, 0, sizeof(RT), s));
      band_queue.alloc(8, s);   // rotating pool: launch n uses counter n % 8 and clears (n + 4) % 8
      CUDA_TRY(cudaMemsetAsync(band_queue.data(), 0, 32, s));
      queue_uses = 0;
      band_launches = 0;
      CUDA_TRY(cudaFuncSetAttribute(sell_band_kernel<RT, true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    static_cast<int>(sell_geom<RT>::smem)));
      CUDA_TRY(cudaFuncSetAttribute(sell_band_kernel<RT, false>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    static_cast<int>(sell_geom<RT>::smem)));
    }
  }
  if (!sell && tiled && !is_fused && !ep->offsets64 && env_int("B200_BANDED", kBandedDefault) != 0) {
    const int T0 = T, rowcap0 = rowcap;
    if (cfg_id == 0 && env_int("B200_BANDED_CFG", 1) != 0) {
      banded_tiles = true;
      if (weighted) {
        T      = banded_cfg<RT, true>::type::T;
        rowcap = banded_cfg<RT, true>::type::ROWCAP;
      } else {
        T      = banded_cfg<RT, false>::type::T;
        rowcap = banded_cfg<RT, false>::type::ROWCAP;
      }
    }
    banded_layout* bl = banded_get<RT>(h, *ep, weights, T, rowcap);
    if (bl && bl->enabled) {
      banded   = bl;
      kp       = bl->reduced.get();
      kweights = weighted ? reinterpret_cast<const RT*>(kp->weights.data()) : nullptr;
      pval.alloc(static_cast<size_t>(bl->num_partials), s, 8);
      // the scratch-column kernel wins only for unweighted 4-byte values (profiles/r1m/)
      band_kernel = env_int("B200_BAND_KERNEL", (sizeof(RT) == 4 && !weighted) ? 2 : 1) == 1 ? 1 : 2;
      band_queue.alloc(2, s);
      CUDA_TRY(cudaMemsetAsync(band_queue.data(), 0, 8, s));
      band_launches = 0;
      // the attribute is per device: one flag per device ordinal (a second handle on another
      // GPU of the process must opt in as well)
      static std::array<std::atomic<bool>, 64> configured{};
      if (!configured[h->device & 63].exchange(true)) {
        CUDA_TRY(cudaFuncSetAttribute(band_reduce_kernel<RT, true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      band_geom<RT>::entries * static_cast<int>(sizeof(RT))));
        CUDA_TRY(cudaFuncSetAttribute(band_reduce_kernel<RT, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      band_geom<RT>::entries * static_cast<int>(sizeof(RT))));
        CUDA_TRY(cudaFuncSetAttribute(band_reduce2_kernel<RT, true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(band_geom2<RT>::smem)));
        CUDA_TRY(cudaFuncSetAttribute(band_reduce2_kernel<RT, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(band_geom2<RT>::smem)));
      }
    } else {
      banded_tiles = false;
      T            = T0;
      rowcap       = rowcap0;
    }
  }
  if (sell) {
    sell_ctas1 = h->sm_count;
    int per_sm = std::max(1, env_int("B200_SELL_CTAS_PER_SM", 8));
    if (g->multi_gpu && h->size > 1) {
      // per-partition launches overlap the previous partition's reduction (NCCL needs SMs of its
      // own: these kernels fill the register file); the fused rectangle overlaps only the deferred
      // 2-scalar allreduce of the previous iteration, and nothing at all with the peer-memory
      // exchange (no NCCL kernel is left in the iteration)
      sell_ctas1 = std::max(1, sell_ctas1 - std::max(0, env_int("B200_MG_RESERVED_SMS", is_fused ? (env_int("B200_MG_P2P", 1) != 0 ? 0 : 4) : 16)));
    }
    main_ctas = sell_ctas1 * per_sm;   // phase-2 grid: 256-thread CTAs, tasks claimed dynamically
    int64_t want_tail = div_up(std::max<int64_t>(sell->num_long_rows, 1), 32);
    tail_ctas         = static_cast<int>(std::min<int64_t>(want_tail, 8 * h->sm_count));
    slots = main_ctas + tail_ctas;
    chunk_partials.alloc(static_cast<size_t>(std::max<int64_t>(sell->num_chunks, 1)), s);
    cta_partials.alloc(static_cast<size_t>(2 * slots), s);
    CUDA_TRY(cudaMemsetAsync(cta_partials.data(), 0, sizeof(double) * 2 * slots, s));
    return;
  }
  if (tiled) {
    if (!banded) build_schedule(h, *ep, T, rowcap);
    main_ctas = h->sm_count;  // one persistent warp-specialised CTA per SM
    // A CTA of this kernel takes the whole register file of its SM, so NCCL's kernels cannot
    // share an SM with it.  On a multi-GPU grid the per-partition reductions run on a second
    // stream while the next partition's kernel executes: leave them a few SMs.
    if (g->multi_gpu && h->size > 1) {
      // the fused rectangle overlaps only the 2-scalar allreduce of the previous iteration
      const bool fused = !g->fused.empty() && ep == &g->fused[0];
      main_ctas = std::max(1, main_ctas - std::max(0, env_int("B200_MG_RESERVED_SMS", fused ? 4 : 16)));
    }
    if (kp->num_tiles < main_ctas) main_ctas = static_cast<int>(std::max<int64_t>(kp->num_tiles, 1));
    // tail kernel: 8 lanes per long row, 32 rows per 256-thread block
    int64_t want_tail = div_up(std::max<int64_t>(kp->num_long_rows, 1), 32);
    tail_ctas         = static_cast<int>(std::min<int64_t>(want_tail, 8 * h->sm_count));
  } else {
    main_ctas = h->sm_count * 8;
    tail_ctas = 1;
  }
  slots = main_ctas + tail_ctas;
  chunk_partials.alloc(static_cast<size_t>(std::max<int64_t>(kp->num_chunks, 1)), s);
  cta_partials.alloc(static_cast<size_t>(2 * slots), s);
  CUDA_TRY(cudaMemsetAsync(cta_partials.data(), 0, sizeof(double) * 2 * slots, s));
}

template <typename RT>
template <typename EPI>
void pull_launcher<RT>::spmv(const RT* x, EPI const& epi, const iter_state* state,
                             bool want_partials)
{
  cudaStream_t s = h->stream;
  double* parts  = want_partials ? cta_partials.data() : nullptr;
  if (!sell && split_event) {   // layouts without a splittable first phase: just wait
    CUDA_TRY(cudaStreamWaitEvent(s, split_event, 0));
    split_event = nullptr;
  }
  if (sell) {
    sell_params1<RT> A{};
    A.x          = x;
    A.pval       = pval.data();
    A.idx        = sell->p1_idx.data();
    A.w          = weighted ? reinterpret_cast<const RT*>(sell->p1_w.data()) : nullptr;
    A.sbase      = sell->p1_sbase.data();
    A.slot       = sell->p1_slot.data();
    A.tasks      = sell->p1_tasks.data();
    A.units      = sell->p1_units.data();
    A.num_minors = ep->num_minors;
    A.state      = state;
    const size_t smem = sell_geom<RT>::smem;
    auto phase1 = [&](int64_t u0, int64_t u1) {
      if (u1 <= u0) return;
      A.unit_first = static_cast<int32_t>(u0);
      A.unit_end   = static_cast<int32_t>(u1);
      A.queue      = band_queue.data() + (queue_uses & 7);
      A.queue_next = band_queue.data() + ((queue_uses + 4) & 7);
      ++queue_uses;
      const int grid = static_cast<int>(std::min<int64_t>(sell_ctas1, u1 - u0));
      if (weighted) sell_band_kernel<RT, true><<<grid, kBandThreads, smem, s>>>(A);
      else          sell_band_kernel<RT, false><<<grid, kBandThreads, smem, s>>>(A);
      CUDA_TRY(cudaGetLastError());
      ++launches;
    };
    const int64_t nu = sell->p1_num_units;
    if (split_event) {
      // units are in band order: those whose band lies inside [split_lo, split_hi) first, the
      // others once the rest of x has arrived
      const int64_t B = sell->band_entries;
      auto const& ub  = sell->p1_unit_band;
      int64_t u0 = 0, u1 = 0;
      while (u0 < nu && static_cast<int64_t>(ub[u0]) * B < split_lo) ++u0;
      u1 = u0;
      while (u1 < nu && (static_cast<int64_t>(ub[u1]) + 1) * B <= split_hi) ++u1;
      phase1(u0, u1);
      CUDA_TRY(cudaStreamWaitEvent(s, split_event, 0));
      phase1(0, u0);
      phase1(u1, nu);
      split_event = nullptr;
    } else {
      phase1(0, nu);
    }
    sell_params2<RT> Q{};
    Q.x              = x;
    Q.pval           = pval.data();
    Q.x_split        = sell->x_split;
    Q.idx            = sell->p2_idx.data();
    Q.w              = weighted ? reinterpret_cast<const RT*>(sell->p2_w.data()) : nullptr;
    Q.sbase          = sell->p2_sbase.data();
    Q.row            = sell->p2_row.data();
    Q.chunk_slices   = sell->p2_chunk_slices;
    Q.num_chunks     = sell->num_chunks;
    Q.task_first     = sell->p2_task_first.data();
    Q.num_tasks      = static_cast<int32_t>(sell->p2_num_tasks);
    Q.queue          = band_queue.data() + (queue_uses & 7);
    Q.queue_next     = band_queue.data() + ((queue_uses + 4) & 7);
    ++queue_uses;
    Q.chunk_partials = chunk_partials.data();
    Q.cta_partials   = parts;
    Q.slots          = slots;
    Q.state          = state;
    Q.hot_limit      = static_cast<int32_t>(std::min<int64_t>(
      ep->num_minors, (static_cast<int64_t>(env_int("B200_HOT_MB", 40)) << 20) / static_cast<int64_t>(sizeof(RT))));
    // L2 policy hints + streaming epilogue pay when the gathered vector and the partial sums are
    // far beyond L2 (scale 26 f64, 1.3 GB: +25 %); up to a few times L2 plain loads are 2-3 %
    // faster (scale 24 f32, 137 MB: 493 vs 481 GTEPS; scale 24 f64, 302 MB: 370 vs 362;
    // profiles/r2_notes.md).  B200_SELL_HINTS overrides.
    {
      const int64_t ws = (ep->num_minors + sell->num_partials) * static_cast<int64_t>(sizeof(RT));
      Q.use_hints      = env_int("B200_SELL_HINTS", ws > (int64_t{512} << 20) ? 1 : 0);
    }
    ++band_launches;
    if (weighted) sell_rows_kernel<RT, true, EPI><<<main_ctas, kSell2Threads, 0, s>>>(Q, epi);
    else          sell_rows_kernel<RT, false, EPI><<<main_ctas, kSell2Threads, 0, s>>>(Q, epi);
    CUDA_TRY(cudaGetLastError());
    ++launches;
    return;
  }
  if (!tiled) {
    auto go = [&](auto et, auto wtag) {
      using ET            = decltype(et);
      constexpr bool W    = decltype(wtag)::value;
      pull_rows_generic_kernel<ET, RT, W, EPI><<<main_ctas, 256, 0, s>>>(
        reinterpret_cast<const ET*>(ep->offsets.data()), ep->indices.data(), weights, x,
        ep->num_rows, epi, state, parts, slots);
    };
    if (ep->offsets64) {
      if (weighted) go(int64_t{}, std::true_type{}); else go(int64_t{}, std::false_type{});
    } else {
      if (weighted) go(int32_t{}, std::true_type{}); else go(int32_t{}, std::false_type{});
    }
    CUDA_TRY(cudaGetLastError());
    ++launches;
    return;
  }
  if (banded) {
    // phase 1: pre-reduce the cells of the hot bands from shared memory
    band_params B{};
    B.x               = x;
    B.pval            = pval.data();
    B.bidx            = banded->bidx.data();
    B.bw              = weighted ? banded->bw.data() : nullptr;
    B.ppos            = banded->ppos.data();
    B.tasks           = banded->tasks.data();
    B.groups          = banded->groups.data();
    B.num_groups      = static_cast<int32_t>(banded->num_groups);
    B.queue           = band_queue.data() + (band_launches & 1);
    B.queue_next      = band_queue.data() + ((band_launches + 1) & 1);
    ++band_launches;
    B.num_minors      = ep->num_minors;
    B.state           = state;
    if (band_kernel == 2) {
      const size_t smem = band_geom2<RT>::smem;
      if (weighted)
        band_reduce2_kernel<RT, true><<<main_ctas, kBandThreads, smem, s>>>(B);
      else
        band_reduce2_kernel<RT, false><<<main_ctas, kBandThreads, smem, s>>>(B);
    } else {
      const size_t smem = static_cast<size_t>(band_geom<RT>::entries) * sizeof(RT);
      if (weighted)
        band_reduce_kernel<RT, true><<<main_ctas, kBandThreads, smem, s>>>(B);
      else
        band_reduce_kernel<RT, false><<<main_ctas, kBandThreads, smem, s>>>(B);
    }
    CUDA_TRY(cudaGetLastError());
    ++launches;
  }
  pull_params P{};
  P.tiles          = kp->tiles.data();
  P.num_tiles      = kp->num_tiles;
  P.offsets        = kp->offsets.data();
  P.indices        = kp->indices.data();
  P.weights        = kweights;
  P.x              = x;
  P.x_hi           = nullptr;
  P.x_split        = INT_MAX;
  P.zero_first     = kp->zero_first;
  P.num_rows       = kp->num_rows;
  P.chunk_partials = chunk_partials.data();
  P.cta_partials   = parts;
  P.slots          = slots;
  P.state          = state;
  if (banded) {
    P.x_hi    = pval.data();
    P.x_split = banded->x_split;
  }
  const int hub_cap = env_int("B200_HUB_ENTRIES", 1 << 30);
  auto launch = [&](auto et, auto wtag, auto cfgtag) {
    using ET         = decltype(et);
    constexpr bool W = decltype(wtag)::value;
    using CFG        = typename decltype(cfgtag)::type;
    using SM         = pull_smem<ET, RT, W, CFG::T, CFG::STAGES, CFG::ROWCAP, EPI::kStage>;
    static_assert(SM::ring <= SM::max_dyn, "tile configuration exceeds the shared memory of an SM");
    auto kern        = pull_tiles_kernel<ET, RT, W, EPI, CFG::T, CFG::STAGES, CFG::ROWCAP>;
    // the hub cache takes what is left of the shared-memory budget; the budget leaves L1 enough
    // lines to track the outstanding gathers (profiles/microbench_gather.md)
    const int64_t budget = static_cast<int64_t>(env_int("B200_SMEM_BUDGET_KB", 160)) * 1024;
    int64_t room = std::max<int64_t>(0, budget - static_cast<int64_t>(SM::ring)) / sizeof(RT);
    int hub = static_cast<int>(std::min<int64_t>({kp->num_minors, SM::max_hub, room})) & ~31;
    hub     = std::max(0, std::min(hub, hub_cap));
    // the entries phase 2 still gathers from x are spread over millions of sources (5 % hit a
    // 16 K cache): no hub cache, the shared memory goes to L1 instead
    if (banded && env_int("B200_BANDED_HUB", 0) == 0) hub = 0;
    static std::array<std::atomic<bool>, 64> configured{};   // per device ordinal
    if (!configured[h->device & 63].exchange(true)) {
      CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    static_cast<int>(SM::total(SM::max_hub))));
    }
    // split the cache between the pieces of the minor range (one piece on a single GPU)
    std::vector<int64_t> pieces = kp->minor_piece_offsets;
    if (pieces.size() < 2 || pieces.size() - 1 > kMaxHubPieces) pieces = {0, kp->num_minors};
    const int np  = static_cast<int>(pieces.size()) - 1;
    int total     = 0;
    P.hub_pieces  = np;
    for (int k = 0; k < kMaxHubPieces; ++k) { P.hub_piece_start[k] = 0; P.hub_piece_len[k] = 0; }
    for (int k = 0; k < np; ++k) {
      int64_t len = std::min<int64_t>((hub / np) & ~31, pieces[k + 1] - pieces[k]);
      P.hub_piece_start[k] = static_cast<int32_t>(pieces[k]);
      P.hub_piece_len[k]   = static_cast<int32_t>(len);
      total += static_cast<int>(len);
    }
    P.hub_entries = total;
    kern<<<main_ctas, kPullThreads, SM::total(hub), s>>>(P, epi);
  };
  auto go = [&](auto et, auto wtag) {
    using ET         = decltype(et);
    constexpr bool W = decltype(wtag)::value;
    if constexpr (sizeof(RT) == 4 && !W && std::is_same<ET, int32_t>::value) {
      switch (cfg_id) {
        case 1: launch(et, wtag, alt_cfg<1>{}); return;
        case 2: launch(et, wtag, alt_cfg<2>{}); return;
        case 3: launch(et, wtag, alt_cfg<3>{}); return;
        case 4: launch(et, wtag, alt_cfg<4>{}); return;
        case 5: launch(et, wtag, alt_cfg<5>{}); return;
        case 6: launch(et, wtag, alt_cfg<6>{}); return;
        case 7: launch(et, wtag, alt_cfg<7>{}); return;
        case 8: launch(et, wtag, alt_cfg<8>{}); return;
        case 9: launch(et, wtag, alt_cfg<9>{}); return;
        case 10: launch(et, wtag, alt_cfg<10>{}); return;
        default: break;
      }
    }
    if (banded_tiles) launch(et, wtag, banded_cfg<RT, W>{});
    else launch(et, wtag, default_cfg<RT, W>{});
  };
  if (kp->offsets64) {
    if (weighted) go(int64_t{}, std::true_type{}); else go(int64_t{}, std::false_type{});
  } else {
    if (weighted) go(int32_t{}, std::true_type{}); else go(int32_t{}, std::false_type{});
  }
  CUDA_TRY(cudaGetLastError());
  ++launches;
}

template <typename RT>
template <typename EPI>
void pull_launcher<RT>::tail(EPI const& epi, iter_state* state, double epsilon, int64_t iteration,
                             bool check, bool finalize)
{
  tail_params P{};
  P.long_row_ids         = (tiled && kp->generic_schedule) ? kp->long_row_ids.data() : nullptr;
  P.long_row_chunk_first = tiled ? kp->long_row_chunk_first.data() : nullptr;
  P.chunk_partials       = chunk_partials.data();
  P.num_long_rows        = tiled ? kp->num_long_rows : 0;
  if (sell) {
    P.long_row_ids         = sell->long_row_ids.data();
    P.long_row_chunk_first = sell->long_row_chunk_first.data();
    P.num_long_rows        = sell->num_long_rows;
  }
  P.cta_partials         = finalize ? cta_partials.data() : nullptr;
  P.slots                = slots;
  P.main_ctas            = main_ctas;
  P.state                = state;
  P.epsilon              = epsilon;
  P.iteration            = iteration;
  P.check_convergence    = check ? 1 : 0;
  P.finalize             = finalize ? 1 : 0;
  if (!finalize && P.num_long_rows == 0) return;
  pull_tail_kernel<RT, EPI, 256><<<tail_ctas, 256, 0, h->stream>>>(P, epi);
  CUDA_TRY(cudaGetLastError());
  ++launches;
}

// ------------------------------------------------------------------------------------------
// vertex passes (replace the thrust::transform / transform_reduce_v calls of the drivers)
// ------------------------------------------------------------------------------------------
namespace {

// per-block partial sums -> slots; the finalize kernel folds them in a fixed order
template <typename RT>
__global__ void __launch_bounds__(kVThreads) sum_kernel(const RT* v, size_t n, double* partials)
{
  __shared__ double scratch[32];
  double s = 0.0;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    s += static_cast<double>(v[i]);
  s = block_sum(s, scratch);
  if (threadIdx.x == 0) partials[blockIdx.x] = s;
}

// state->a = sum(pa[0..n)), state->b = sum(pb[0..n)) (either may be nullptr)
__global__ void __launch_bounds__(256) fold_kernel(const double* pa, const double* pb, int n,
                                                   iter_state* st)
{
  __shared__ double scratch[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    if (pa) a += pa[i];
    if (pb) b += pb[i];
  }
  a = block_sum(a, scratch);
  b = block_sum(b, scratch);
  if (threadIdx.x == 0) {
    if (pa) st->a = a;
    if (pb) st->b = b;
  }
}

// scale[v] = alpha / out_weight[v], or -1 when out_weight[v] == 0 (dangling)
template <typename RT, typename WT>
__global__ void __launch_bounds__(kVThreads) pagerank_scale_kernel(const WT* outw, size_t n,
                                                                  RT alpha, RT* scale)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    RT w     = static_cast<RT>(outw[i]);
    scale[i] = (w == RT(0)) ? RT(-1) : alpha / w;
  }
}

// pr = guess / sum (or 1/V); x0 = pr * scale; dangling partials
template <typename RT>
__global__ void __launch_bounds__(kVThreads) pagerank_init_kernel(RT* pr, const RT* scale, RT* x0,
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
      x0[i] = RT(0);
    } else {
      x0[i] = v * sc;
    }
  }
  dang = block_sum(dang, scratch);
  if (threadIdx.x == 0) partials[blockIdx.x] = dang;
}

template <typename RT>
__global__ void __launch_bounds__(kVThreads) scatter_pers_kernel(const int32_t* vertices,
                                                                const RT* values, size_t n,
                                                                RT pers_sum, int64_t first,
                                                                int64_t nv, RT* dense, int* bad)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    int64_t v = static_cast<int64_t>(vertices[i]) - first;
    if (v < 0 || v >= nv) { *bad = 1; continue; }
    // duplicates accumulate, as the reference's per-entry += does (pagerank_impl.cuh:259-276)
    atomicAdd(&dense[v], values[i] / pers_sum);
  }
}

template <typename RT>
__global__ void __launch_bounds__(kVThreads) divide_kernel(RT* v, size_t n, RT d)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    v[i] = v[i] / d;
}

template <typename RT>
__global__ void __launch_bounds__(kVThreads) fill_kernel(RT* v, size_t n, RT x)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    v[i] = x;
}

// eigenvector: c = c / norm, diff partial = |c - old|
template <typename RT>
__global__ void __launch_bounds__(kVThreads) eigen_normalize_kernel(RT* c, const RT* old, size_t n,
                                                                   const iter_state* st,
                                                                   double* partials)
{
  __shared__ double scratch[32];
  if (st->done) return;
  RT norm     = static_cast<RT>(sqrt(st->a));
  double diff = 0.0;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    RT v = c[i] / norm;
    c[i] = v;
    diff += fabs(static_cast<double>(v) - static_cast<double>(old[i]));
  }
  diff = block_sum(diff, scratch);
  if (threadIdx.x == 0) partials[blockIdx.x] = diff;
}

// fold diff partials, stop test, iteration bookkeeping (eigenvector)
__global__ void __launch_bounds__(256) eigen_fold_kernel(const double* partials, int n,
                                                         iter_state* st, double threshold,
                                                         int64_t iteration, int check)
{
  __shared__ double scratch[32];
  if (st->done) return;
  double b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) b += partials[i];
  b = block_sum(b, scratch);
  if (threadIdx.x == 0) {
    st->b          = b;
    st->iterations = iteration;
    if (check && b < threshold) {
      st->converged = 1;
      st->done      = 1;
    }
  }
}

struct host_state {
  iter_state st;
};

void read_state(b200_handle* h, const iter_state* d, iter_state* out)
{
  CUDA_TRY(cudaMemcpyAsync(out, d, sizeof(iter_state), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
}

struct loop_timer {
  cudaEvent_t e0{}, e1{};
  cudaStream_t s;
  explicit loop_timer(cudaStream_t s_) : s(s_)
  {
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
  }
  ~loop_timer()
  {
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
  }
  void start() { cudaEventRecord(e0, s); }
  double stop()
  {
    cudaEventRecord(e1, s);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
  }
};

// CUDA events around every pull-kernel launch (first kMax iterations): the roofline figure of
// bench.py is algorithmic bytes / this duration
struct launch_timer {
  static constexpr int kMax = 256;
  std::vector<cudaEvent_t> ev;
  cudaStream_t s;
  int n{0};
  explicit launch_timer(cudaStream_t s_) : s(s_) {}
  ~launch_timer() { for (auto e : ev) cudaEventDestroy(e); }
  void before()
  {
    if (n >= kMax) return;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    ev.push_back(a);
    ev.push_back(b);
    cudaEventRecord(a, s);
  }
  void after()
  {
    if (n >= kMax) return;
    cudaEventRecord(ev[2 * n + 1], s);
    ++n;
  }
  // call after the stream has been synchronised
  void collect(b200_centrality_stats_t* st)
  {
    if (!st) return;
    double tot = 0.0;
    for (int i = 0; i < n; ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[2 * i], ev[2 * i + 1]);
      tot += ms;
    }
    st->spmv_ms       = tot;
    st->spmv_launches = n;
  }
};

}  // namespace

// ------------------------------------------------------------------------------------------
// PageRank (single GPU)
// ------------------------------------------------------------------------------------------
template <typename RT>
static void pagerank_sg(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, RT* pr,
                        b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const int64_t nv  = g->num_vertices;
  const size_t n    = static_cast<size_t>(nv);
  const RT alpha    = static_cast<RT>(p->alpha);
  const RT epsilon  = static_cast<RT>(p->epsilon);
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  if (stats) *stats = b200_centrality_stats_t{};
  if (nv == 0) return;

  pull_launcher<RT> L;
  L.init(h, g, &g->parts[0]);
  const unsigned vg = vgrid(n, h->sm_count);

  dbuf<iter_state> d_state(1, s);
  CUDA_TRY(cudaMemsetAsync(d_state.data(), 0, sizeof(iter_state), s));
  dbuf<double> vparts(vg, s);

  // out-weight sums -> scale
  // arrays the TMA producer stages by 16-byte pieces are padded past their logical end
  dbuf<RT> scale(n, s, 8);
  dbuf<RT> pr_work(n, s, 8);
  RT* const pr_user = pr;
  if (p->has_initial_guess)
    CUDA_TRY(cudaMemcpyAsync(pr_work.data(), pr_user, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  pr = pr_work.data();
  if (p->precomputed_out_weight_sums) {
    pagerank_scale_kernel<RT, RT><<<vg, kVThreads, 0, s>>>(
      static_cast<const RT*>(p->precomputed_out_weight_sums), n, alpha, scale.data());
  } else {
    graph_out_weight_sums(h, g);
    // the reference holds the sums in weight_t (graph_view_impl.cuh:238-265); round the same way
    pagerank_scale_kernel<RT, double><<<vg, kVThreads, 0, s>>>(g->out_weight_sums.data(), n, alpha,
                                                              scale.data());
  }
  CUDA_TRY(cudaGetLastError());

  // personalization (dense, normalised)
  dbuf<RT> pers;
  const bool personalized = p->personalization_vertices != nullptr && p->personalization_size > 0;
  if (p->personalization_vertices != nullptr) {
    B200_EXPECTS(p->personalization_values != nullptr, kInvalidInput,
                 "Invalid input argument: if personalization_vertices.has_value() is true, "
                 "personalization_values.has_value() and personalization_vector_size.has_value() "
                 "should be true as well.");
  }
  if (personalized) {
    size_t np   = static_cast<size_t>(p->personalization_size);
    unsigned pg = vgrid(np, h->sm_count);
    dbuf<double> pparts(pg, s);
    sum_kernel<RT><<<pg, kVThreads, 0, s>>>(static_cast<const RT*>(p->personalization_values), np,
                                            pparts.data());
    fold_kernel<<<1, 256, 0, s>>>(pparts.data(), nullptr, static_cast<int>(pg), d_state.data());
    CUDA_TRY(cudaGetLastError());
    iter_state hs;
    read_state(h, d_state.data(), &hs);
    RT pers_sum = static_cast<RT>(hs.a);
    B200_EXPECTS(pers_sum > RT(0), kInvalidInput,
                 "Invalid input argument: sum of personalization valuese should be positive.");
    pers.alloc(n, s);
    CUDA_TRY(cudaMemsetAsync(pers.data(), 0, n * sizeof(RT), s));
    dbuf<int> bad(1, s);
    CUDA_TRY(cudaMemsetAsync(bad.data(), 0, 4, s));
    scatter_pers_kernel<RT><<<pg, kVThreads, 0, s>>>(
      static_cast<const int32_t*>(p->personalization_vertices),
      static_cast<const RT*>(p->personalization_values), np, pers_sum, 0, nv, pers.data(),
      bad.data());
    CUDA_TRY(cudaGetLastError());
    int h_bad = 0;
    CUDA_TRY(cudaMemcpyAsync(&h_bad, bad.data(), 4, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    B200_EXPECTS(!h_bad, kInvalidInput,
                 "Invalid input argument: peresonalization vertices have invalid vertex IDs.");
  }

  // initial values
  if (p->has_initial_guess) {
    sum_kernel<RT><<<vg, kVThreads, 0, s>>>(pr, n, vparts.data());
    fold_kernel<<<1, 256, 0, s>>>(vparts.data(), nullptr, static_cast<int>(vg), d_state.data());
    CUDA_TRY(cudaGetLastError());
    iter_state hs;
    read_state(h, d_state.data(), &hs);
    B200_EXPECTS(static_cast<RT>(hs.a) > RT(0), kInvalidInput,
                 "Invalid input argument: sum of the PageRank initial guess values should be "
                 "positive.");
  }
  dbuf<RT> xbuf(2 * n, s);
  RT* x[2] = {xbuf.data(), xbuf.data() + n};
  pagerank_init_kernel<RT><<<vg, kVThreads, 0, s>>>(pr, scale.data(), x[0], n,
                                                    p->has_initial_guess ? 1 : 0, d_state.data(),
                                                    RT(1) / static_cast<RT>(nv), vparts.data());
  fold_kernel<<<1, 256, 0, s>>>(vparts.data(), nullptr, static_cast<int>(vg), d_state.data());
  CUDA_TRY(cudaGetLastError());

  epi_pagerank<RT> epi{};
  epi.pr                  = pr;
  epi.scale               = scale.data();
  epi.pers                = personalized ? pers.data() : nullptr;
  epi.alpha               = alpha;
  epi.one_minus_alpha     = static_cast<RT>(1.0 - p->alpha);
  epi.num_vertices        = static_cast<RT>(nv);

  const int check_every = std::max(1, env_int("B200_CHECK_EVERY", 8));
  loop_timer timer(s);
  launch_timer ktimer(s);
  timer.start();
  iter_state hs{};
  int64_t it = 0;
  for (; it < cap; ++it) {
    epi.x_next = x[(it + 1) & 1];
    ktimer.before();
    L.spmv(x[it & 1], epi, d_state.data(), true);
    ktimer.after();
    L.tail(epi, d_state.data(), static_cast<double>(epsilon), it + 1, !fixed, true);
    if (!fixed && ((it + 1) % check_every == 0)) {
      read_state(h, d_state.data(), &hs);
      if (hs.done) break;
    }
  }
  double ms = timer.stop();
  read_state(h, d_state.data(), &hs);
  if (stats) {
    stats->iterations   = hs.iterations;
    stats->converged    = hs.converged;
    stats->last_diff    = hs.b;
    stats->elapsed_ms   = ms;
    stats->gpu_launches = L.launches;
    ktimer.collect(stats);
  }
  CUDA_TRY(cudaMemcpyAsync(pr_user, pr, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  if (!fixed && !hs.converged) {
    throw error(kUnknownError, "PageRank failed to converge.");
  }
}

void pagerank(b200_handle* h, b200_graph* g, const b200_pagerank_params_t* p, void* pageranks,
              b200_centrality_stats_t* stats)
{
  B200_EXPECTS(g->store_transposed, kInvalidInput,
               "PageRank needs a graph stored transposed (pull layout).");
  B200_EXPECTS((p->alpha >= 0.0) && (p->alpha <= 1.0), kInvalidInput,
               "Invalid input argument: alpha should be in [0.0, 1.0].");
  B200_EXPECTS(p->epsilon >= 0.0, kInvalidInput,
               "Invalid input argument: epsilon should be non-negative.");
  B200_EXPECTS(p->value_type == B200_FLOAT32 || p->value_type == B200_FLOAT64,
               kUnsupportedTypeCombination, "result type must be FLOAT32 or FLOAT64");
  if (g->multi_gpu) {
    pagerank_mg(h, g, p, pageranks, stats);
    return;
  }
  if (p->value_type == B200_FLOAT32)
    pagerank_sg<float>(h, g, p, static_cast<float*>(pageranks), stats);
  else
    pagerank_sg<double>(h, g, p, static_cast<double*>(pageranks), stats);
}

// ------------------------------------------------------------------------------------------
// Katz centrality (single GPU)
// ------------------------------------------------------------------------------------------
template <typename RT>
static void katz_sg(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, RT* kc,
                    b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const int64_t nv  = g->num_vertices;
  const size_t n    = static_cast<size_t>(nv);
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  const RT epsilon  = static_cast<RT>(p->epsilon);
  if (stats) *stats = b200_centrality_stats_t{};
  if (nv == 0) return;

  pull_launcher<RT> L;
  L.init(h, g, &g->parts[0]);
  const unsigned vg = vgrid(n, h->sm_count);
  dbuf<iter_state> d_state(1, s);
  CUDA_TRY(cudaMemsetAsync(d_state.data(), 0, sizeof(iter_state), s));
  dbuf<RT> work0(n, s, 8), work1(n, s, 8);  // padded: staged by 16-byte pieces
  if (p->has_initial_guess)
    CUDA_TRY(cudaMemcpyAsync(work0.data(), kc, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  else
    CUDA_TRY(cudaMemsetAsync(work0.data(), 0, n * sizeof(RT), s));
  // buffers alternate; iteration `it` reads buf[it & 1] and writes buf[(it + 1) & 1]
  RT* buf[2] = {work0.data(), work1.data()};

  epi_katz<RT> epi{};
  epi.betas = static_cast<const RT*>(p->betas);
  epi.alpha = static_cast<RT>(p->alpha);
  epi.beta  = static_cast<RT>(p->beta);

  const int check_every = std::max(1, env_int("B200_CHECK_EVERY", 8));
  loop_timer timer(s);
  timer.start();
  iter_state hs{};
  for (int64_t it = 0; it < cap; ++it) {
    epi.old_c = buf[it & 1];
    epi.new_c = buf[(it + 1) & 1];
    L.spmv(epi.old_c, epi, d_state.data(), true);
    L.tail(epi, d_state.data(), static_cast<double>(epsilon), it + 1, !fixed, true);
    if (!fixed && ((it + 1) % check_every == 0)) {
      read_state(h, d_state.data(), &hs);
      if (hs.done) break;
    }
  }
  double ms = timer.stop();
  read_state(h, d_state.data(), &hs);
  if (stats) {
    stats->iterations   = hs.iterations;
    stats->converged    = hs.converged;
    stats->last_diff    = hs.b;
    stats->elapsed_ms   = ms;
    stats->gpu_launches = L.launches;
  }
  if (!fixed && !hs.converged) throw error(kUnknownError, "Katz Centrality failed to converge.");
  RT* result = buf[hs.iterations & 1];
  CUDA_TRY(cudaMemcpyAsync(kc, result, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  if (p->normalize) {
    RT l2 = static_cast<RT>(std::sqrt(static_cast<RT>(hs.a)));
    B200_EXPECTS(l2 > RT(0), kUnknownError,
                 "L2 norm of the computed Katz Centrality values should be positive.");
    divide_kernel<RT><<<vg, kVThreads, 0, s>>>(kc, n, l2);
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaStreamSynchronize(s));
}

void katz_centrality(b200_handle* h, b200_graph* g, const b200_katz_params_t* p, void* out,
                     b200_centrality_stats_t* stats)
{
  B200_EXPECTS(g->store_transposed, kInvalidInput,
               "Katz centrality needs a graph stored transposed (pull layout).");
  B200_EXPECTS((p->alpha >= 0.0) && (p->alpha <= 1.0), kInvalidInput,
               "Invalid input argument: alpha should be in [0.0, 1.0].");
  B200_EXPECTS(p->epsilon >= 0.0, kInvalidInput,
               "Invalid input argument: epsilon should be non-negative.");
  if (g->multi_gpu) {
    katz_mg(h, g, p, out, stats);
    return;
  }
  if (p->value_type == B200_FLOAT32) katz_sg<float>(h, g, p, static_cast<float*>(out), stats);
  else katz_sg<double>(h, g, p, static_cast<double*>(out), stats);
}

// ------------------------------------------------------------------------------------------
// eigenvector centrality (single GPU)
// ------------------------------------------------------------------------------------------
template <typename RT>
static void eigen_sg(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p, RT* c,
                     b200_centrality_stats_t* stats)
{
  cudaStream_t s    = h->stream;
  const int64_t nv  = g->num_vertices;
  const size_t n    = static_cast<size_t>(nv);
  const bool fixed  = p->fixed_iterations > 0;
  const int64_t cap = fixed ? p->fixed_iterations
                            : std::max<int64_t>(1, static_cast<int64_t>(p->max_iterations));
  if (stats) *stats = b200_centrality_stats_t{};
  if (nv == 0) return;
  pull_launcher<RT> L;
  L.init(h, g, &g->parts[0]);
  const unsigned vg = vgrid(n, h->sm_count);
  dbuf<iter_state> d_state(1, s);
  CUDA_TRY(cudaMemsetAsync(d_state.data(), 0, sizeof(iter_state), s));
  dbuf<double> vparts(vg, s);
  dbuf<RT> tmp(n, s);
  if (!p->has_initial_guess) {
    fill_kernel<RT><<<vg, kVThreads, 0, s>>>(c, n, RT(1) / static_cast<RT>(nv));
    CUDA_TRY(cudaGetLastError());
  }
  RT* buf[2] = {c, tmp.data()};
  epi_eigen<RT> epi{};
  // the reference compares against number_of_vertices * epsilon in weight_t arithmetic
  const double threshold = static_cast<double>(static_cast<RT>(nv) * static_cast<RT>(p->epsilon));
  const int check_every  = std::max(1, env_int("B200_CHECK_EVERY", 8));
  loop_timer timer(s);
  timer.start();
  iter_state hs{};
  int64_t extra = 0;
  for (int64_t it = 0; it < cap; ++it) {
    RT* old_c = buf[it & 1];
    epi.new_c = buf[(it + 1) & 1];
    L.spmv(old_c, epi, d_state.data(), true);
    // tail folds sum of squares into state.a (no stop test here)
    L.tail(epi, d_state.data(), 0.0, it, false, true);
    eigen_normalize_kernel<RT><<<vg, kVThreads, 0, s>>>(epi.new_c, old_c, n, d_state.data(),
                                                        vparts.data());
    eigen_fold_kernel<<<1, 256, 0, s>>>(vparts.data(), static_cast<int>(vg), d_state.data(),
                                        threshold, it + 1, fixed ? 0 : 1);
    CUDA_TRY(cudaGetLastError());
    extra += 2;
    if (!fixed && ((it + 1) % check_every == 0)) {
      read_state(h, d_state.data(), &hs);
      if (hs.done) break;
    }
  }
  double ms = timer.stop();
  read_state(h, d_state.data(), &hs);
  if (stats) {
    stats->iterations   = hs.iterations;
    stats->converged    = hs.converged;
    stats->last_diff    = hs.b;
    stats->elapsed_ms   = ms;
    stats->gpu_launches = L.launches + extra;
  }
  if (!fixed && !hs.converged)
    throw error(kUnknownError, "Eigenvector Centrality failed to converge.");
  RT* result = buf[hs.iterations & 1];
  if (result != c) CUDA_TRY(cudaMemcpyAsync(c, result, n * sizeof(RT), cudaMemcpyDeviceToDevice, s));
  CUDA_TRY(cudaStreamSynchronize(s));
}

void eigenvector_centrality(b200_handle* h, b200_graph* g, const b200_eigenvector_params_t* p,
                            void* out, b200_centrality_stats_t* stats)
{
  B200_EXPECTS(g->store_transposed, kInvalidInput,
               "Eigenvector centrality needs a graph stored transposed (pull layout).");
  B200_EXPECTS(p->epsilon >= 0.0, kInvalidInput,
               "Invalid input argument: epsilon should be non-negative.");
  if (g->multi_gpu) {
    eigenvector_mg(h, g, p, out, stats);
    return;
  }
  if (p->value_type == B200_FLOAT32) eigen_sg<float>(h, g, p, static_cast<float*>(out), stats);
  else eigen_sg<double>(h, g, p, static_cast<double*>(out), stats);
}

// ------------------------------------------------------------------------------------------
// one edge partition of a multi-GPU graph: y = A_i^T x with the raw epilogue (mg.cu sums the
// partitions' results over col_comm and runs the driver's vertex pass afterwards)
// ------------------------------------------------------------------------------------------
struct partition_spmv {
  int value_type{B200_FLOAT32};
  pull_launcher<float> lf;
  pull_launcher<double> ld;
};

partition_spmv* partition_spmv_create(b200_handle* h, b200_graph* g, int part, int value_type)
{
  auto* p       = new partition_spmv();
  p->value_type = value_type;
  // part < 0: the fused rectangle of a multi-GPU graph (all local partitions in one launch)
  edge_partition* ep = part < 0 ? &g->fused.at(0) : &g->parts.at(part);
  try {
    if (value_type == B200_FLOAT32) p->lf.init(h, g, ep);
    else p->ld.init(h, g, ep);
  } catch (...) {
    delete p;
    throw;
  }
  return p;
}

void partition_spmv_run(partition_spmv* p, const void* x, void* y, const iter_state* state)
{
  if (p->value_type == B200_FLOAT32) {
    if (p->lf.ep->num_rows == 0) return;
    epi_raw<float> epi{static_cast<float*>(y), 1.0f, 0.0f};
    if (p->lf.sell && p->lf.sell->skips_empty_rows)
      CUDA_TRY(cudaMemsetAsync(y, 0, static_cast<size_t>(p->lf.ep->num_rows) * sizeof(float), p->lf.h->stream));
    p->lf.spmv(static_cast<const float*>(x), epi, state, false);
    p->lf.tail(epi, const_cast<iter_state*>(state), 0.0, 0, false, false);
  } else {
    if (p->ld.ep->num_rows == 0) return;
    epi_raw<double> epi{static_cast<double*>(y), 1.0, 0.0};
    if (p->ld.sell && p->ld.sell->skips_empty_rows)
      CUDA_TRY(cudaMemsetAsync(y, 0, static_cast<size_t>(p->ld.ep->num_rows) * sizeof(double), p->ld.h->stream));
    p->ld.spmv(static_cast<const double*>(x), epi, state, false);
    p->ld.tail(epi, const_cast<iter_state*>(state), 0.0, 0, false, false);
  }
}

void partition_spmv_split(partition_spmv* p, int64_t lo, int64_t hi, cudaEvent_t rest_ready)
{
  if (p->value_type == B200_FLOAT32) {
    p->lf.split_lo = lo; p->lf.split_hi = hi; p->lf.split_event = rest_ready;
  } else {
    p->ld.split_lo = lo; p->ld.split_hi = hi; p->ld.split_event = rest_ready;
  }
}

int64_t partition_spmv_launches(const partition_spmv* p)
{
  return p->value_type == B200_FLOAT32 ? p->lf.launches : p->ld.launches;
}

void partition_spmv_free(partition_spmv* p) { delete p; }

// ------------------------------------------------------------------------------------------
// raw SpMV: y = init + scale * A^T x
// ------------------------------------------------------------------------------------------
template <typename RT>
static void spmv_sg(b200_handle* h, b200_graph* g, const RT* x, double scale, double init, RT* y)
{
  if (g->num_vertices == 0) return;
  pull_launcher<RT> L;
  L.init(h, g, &g->parts[0]);
  epi_raw<RT> epi{y, static_cast<RT>(scale), static_cast<RT>(init)};
  L.spmv(x, epi, nullptr, false);
  L.tail(epi, nullptr, 0.0, 0, false, false);
  CUDA_TRY(cudaStreamSynchronize(h->stream));
}

void spmv_pull(b200_handle* h, b200_graph* g, int value_type, const void* x, double scale,
               double init, void* y)
{
  B200_EXPECTS(!g->multi_gpu, kNotImplemented, "b200_spmv_pull is single-GPU");
  if (value_type == B200_FLOAT32)
    spmv_sg<float>(h, g, static_cast<const float*>(x), scale, init, static_cast<float*>(y));
  else
    spmv_sg<double>(h, g, static_cast<const double*>(x), scale, init, static_cast<double*>(y));
}

}  // namespace b200

#ifdef B200_PROFILE_ROLES
// tuning builds only (scripts/build_variant.sh prof -DB200_PROFILE_ROLES)
extern "C" int b200_debug_role_clocks(unsigned long long* out)
{
  return cudaMemcpyFromSymbol(out, b200::b200_role_clocks, sizeof(unsigned long long) * 256 * 8) ==
             cudaSuccess
           ? 0
           : 1;
}
#endif
