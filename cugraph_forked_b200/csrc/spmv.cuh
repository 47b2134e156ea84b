// The hot path: pull-direction edge->vertex transform-reduce over a CSC rectangle, hand-written
// for sm_100a.  B200-native replacement of the four kernels of
// cpp/src/prims/per_v_transform_reduce_incoming_outgoing_e.cuh (:84 hypersparse, :194 low,
// :306 mid, :401 high) and of their host orchestration (:489-1038).
//
// Design (see DESIGN.md "kernels"):
//  * the edge arrays are cut ONCE (at graph build) into tiles of <= T edges that respect the
//    reference's degree segments: rows of the high-degree segment (degree >= 1024,
//    graph_view.hpp:255-263) are split into chunks of T edges that many CTAs reduce
//    cooperatively (block-wide reduction + a deterministic combine in the tail kernel);
//    mid- and low-degree rows are packed into tiles of COMPLETE rows.
//  * a persistent grid (CTAS_PER_SM x 148 CTAs) walks the tiles round-robin.  One elected
//    thread streams each tile's indices / weights / offsets into shared memory with
//    cp.async.bulk (the TMA bulk-copy engine, SASS UBLKCP) into a STAGES-deep ring guarded by
//    mbarriers; index traffic carries an L2 evict-first policy so that it does not displace
//    the gathered vector.
//  * every thread reads 4 indices at a time from shared memory (128-bit LDS), issues the 4
//    independent gathers x[idx], and writes the products back to shared memory.
//  * rows are then reduced from shared memory by groups of G = 1..32 lanes per row
//    (thread-per-vertex for the low-degree segment, sub-warp / warp segmented shuffles for
//    the mid-degree segment), G chosen per tile from its mean degree.
//  * the caller's epilogue (PageRank: + unvarying part, new scaled vector, dangling mass and
//    L1 difference; Katz: + beta, L1 difference; raw: y = init + scale*sum) is fused into the
//    row reduction, so one launch per power iteration does what the reference does in
//    ~10 launches and 2 host synchronisations (pagerank_impl.cuh:209-292).
#pragma once

#include "common.cuh"
#include "prims.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + bulk async copy (TMA 1D) + cache policies
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init()
{
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity)
{
  uint32_t ok;
  asm volatile(
    "{\n"
    ".reg .pred p;\n"
    "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
    "selp.u32 %0, 1, 0, p;\n"
    "}\n"
    : "=r"(ok)
    : "r"(smem_u32(bar)), "r"(parity)
    : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
  while (!mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ uint64_t policy_evict_first()
{
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last()
{
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar, uint64_t policy)
{
  asm volatile(
    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
    "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst_smem)),
    "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
    : "memory");
}

// gather load of the vertex-value vector: read-only path, keep in L1/L2 as long as possible
template <typename T>
__device__ __forceinline__ T ld_gather(const T* p)
{
  return __ldg(p);
}

// ------------------------------------------------------------------------------------------
// per-thread accumulators of the fused epilogues (two scalars per iteration)
// ------------------------------------------------------------------------------------------
struct acc2 {
  double a{0.0};  // PageRank: dangling mass of the NEW vector; Katz/eigen: sum of squares
  double b{0.0};  // L1 difference |new - old|
};

// device-resident iteration state shared by all kernels of a driver loop
struct iter_state {
  double a;          // finalised `a` of the last completed iteration
  double b;          // finalised `b`
  double aux;        // driver-specific (eigenvector: previous norm)
  int64_t iterations;
  int32_t done;
  int32_t converged;
  uint32_t ticket;   // last-block election of the tail kernel
  uint32_t pad;
};

// ---- epilogues ---------------------------------------------------------------------------
// raw: y[row] = init + scale * sum   (b200_spmv_pull, multi-GPU partial sums)
template <typename RT>
struct epi_raw {
  RT* y;
  RT scale;
  RT init;
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2&) const
  {
    y[row] = init + scale * sum;
  }
};

// PageRank (pagerank_impl.cuh:209-292 fused): the gathered vector is
// x[v] = pr[v] * alpha / out_weight[v], so `sum` already is alpha * sum(pr/outw * w).
template <typename RT>
struct epi_pagerank {
  RT* pr;             // in: old, out: new
  RT* x_next;         // next gathered vector
  const RT* scale;    // alpha / out_weight, or -1 for dangling vertices (out_weight == 0)
  const RT* pers;     // personalization value / personalization sum (dense), or nullptr
  RT alpha;
  RT one_minus_alpha;
  RT inv_or_num_vertices;  // static_cast<RT>(num_vertices)
  RT unvarying;       // set by setup()
  RT pers_factor;     // (dangling * alpha + (1 - alpha)) when personalized
  __device__ __forceinline__ void setup(const iter_state* st)
  {
    RT dangling = static_cast<RT>(st->a);
    RT base     = dangling * alpha + one_minus_alpha;
    if (pers) {
      unvarying   = RT(0);
      pers_factor = base;
    } else {
      unvarying   = base / inv_or_num_vertices;
      pers_factor = RT(0);
    }
  }
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    RT old = pr[row];
    RT sc  = scale[row];
    RT v   = sum + unvarying;
    if (pers) v += pers_factor * pers[row];
    pr[row] = v;
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(old));
    if (sc < RT(0)) {
      acc.a += static_cast<double>(v);
      x_next[row] = RT(0);
    } else {
      x_next[row] = v * sc;
    }
  }
};

// Katz (katz_centrality_impl.cuh:104-148): new = alpha * sum + beta (or betas[v]);
// the gathered vector is the old centrality itself.
template <typename RT>
struct epi_katz {
  const RT* old_c;
  RT* new_c;
  const RT* betas;  // or nullptr
  RT alpha;
  RT beta;
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    RT v       = alpha * sum + (betas ? betas[row] : beta);
    new_c[row] = v;
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(old_c[row]));
    acc.a += static_cast<double>(v) * static_cast<double>(v);
  }
};

// eigenvector (eigenvector_centrality_impl.cuh:90-135): new = sum; L2 norm accumulated here,
// normalisation + difference in a separate vertex pass.
template <typename RT>
struct epi_eigen {
  RT* new_c;
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    new_c[row] = sum;
    acc.a += static_cast<double>(sum) * static_cast<double>(sum);
  }
};

// ------------------------------------------------------------------------------------------
// kernel parameters
// ------------------------------------------------------------------------------------------
struct pull_params {
  const tile_t* tiles;
  int64_t num_tiles;
  const void* offsets;     // ET [num_rows + 1]
  const int32_t* indices;  // padded by >= 16 bytes
  const void* weights;     // WT or nullptr
  const void* x;           // gathered vector, RT
  int64_t zero_first;      // rows [zero_first, num_rows) have no edges (epilogue only)
  int64_t num_rows;
  double* chunk_partials;  // [num_chunks]
  double* cta_partials;    // [2 (a,b)][slots]  written at [blockIdx.x]
  int32_t slots;
  const iter_state* state; // may be nullptr (raw spmv)
};

constexpr int kLongRowDegree = 1024;  // = mid_degree_threshold, graph_view.hpp:255-263
constexpr int kRowStageCap   = 1024;  // tiles with more rows read their offsets from global

template <int T>
struct tile_geom {
  // staged element capacity: T edges + one row overhang (< 1024) + alignment slack
  static constexpr int cap = T + kLongRowDegree + 8;
};

template <typename ET, typename WT, typename RT, bool WEIGHTED, int T, int STAGES>
struct pull_smem {
  static constexpr int cap       = tile_geom<T>::cap;
  static constexpr size_t idx_b  = size_t(cap) * 4;
  static constexpr size_t w_b    = WEIGHTED ? size_t(cap) * sizeof(WT) : 0;
  static constexpr size_t off_b  = size_t(kRowStageCap + 8) * sizeof(ET);
  static constexpr size_t stage  = ((idx_b + w_b + off_b) + 127) & ~size_t(127);
  // 4-byte results overwrite the staged indices in place; 8-byte results need their own
  // (double-buffered) product array
  static constexpr bool inplace  = sizeof(RT) == 4;
  static constexpr size_t prod_b = inplace ? 0 : ((size_t(cap) * sizeof(RT) + 127) & ~size_t(127));
  static constexpr size_t total  = STAGES * stage + 2 * prod_b;
};

// ------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------
template <typename ET, typename WT, typename RT, bool WEIGHTED, typename EPI, int T, int THREADS,
          int STAGES, int CTAS_PER_SM>
__global__ void __launch_bounds__(THREADS, CTAS_PER_SM)
  pull_tiles_kernel(const __grid_constant__ pull_params P, EPI epi)
{
  using S = pull_smem<ET, WT, RT, WEIGHTED, T, STAGES>;
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t full_bar[STAGES];
  __shared__ double warp_part[2][THREADS / 32];
  __shared__ double red_scratch[32];

  if (P.state && P.state->done) return;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned char* stage_base = smem;
  RT* prod_base             = reinterpret_cast<RT*>(smem + STAGES * S::stage);

  auto idx_s = [&](int s) { return reinterpret_cast<int32_t*>(stage_base + s * S::stage); };
  auto w_s   = [&](int s) { return reinterpret_cast<WT*>(stage_base + s * S::stage + S::idx_b); };
  auto off_s = [&](int s) {
    return reinterpret_cast<ET*>(stage_base + s * S::stage + S::idx_b + S::w_b);
  };
  auto prod_s = [&](int b) {
    return reinterpret_cast<RT*>(reinterpret_cast<unsigned char*>(prod_base) + b * S::prod_b);
  };

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&full_bar[s], 1);
    fence_mbar_init();
  }
  epi.setup(P.state);
  __syncthreads();

  const ET* offsets    = static_cast<const ET*>(P.offsets);
  const WT* weights    = static_cast<const WT*>(P.weights);
  const RT* x          = static_cast<const RT*>(P.x);
  const int64_t stride = gridDim.x;
  const int64_t my_n =
    (P.num_tiles > blockIdx.x) ? (P.num_tiles - blockIdx.x + stride - 1) / stride : 0;

  uint64_t pol_stream = 0;
  if (tid == 0) pol_stream = policy_evict_first();

  // producer: stream tile j of this CTA into its ring slot
  auto issue = [&](int64_t j) {
    const tile_t t = P.tiles[blockIdx.x + j * stride];
    const int s    = static_cast<int>(j % STAGES);
    const int lead = static_cast<int>(t.edge_first & 3);
    const int64_t a0 = t.edge_first - lead;
    const uint32_t cnt = static_cast<uint32_t>((lead + t.num_edges + 3) & ~3);
    uint32_t bytes     = cnt * 4u;
    if (WEIGHTED) bytes += cnt * static_cast<uint32_t>(sizeof(WT));
    uint32_t off_bytes = 0;
    int64_t r0         = 0;
    if (t.kind == 0 && t.num_rows <= kRowStageCap) {
      constexpr int per16 = 16 / sizeof(ET);
      r0                  = t.row_first & ~int64_t(per16 - 1);
      int n_off           = static_cast<int>(t.row_first - r0) + t.num_rows + 1;
      n_off               = (n_off + per16 - 1) & ~(per16 - 1);
      off_bytes           = static_cast<uint32_t>(n_off * sizeof(ET));
    }
    mbar_arrive_expect_tx(&full_bar[s], bytes + off_bytes);
    bulk_g2s(idx_s(s), P.indices + a0, cnt * 4u, &full_bar[s], pol_stream);
    if (WEIGHTED)
      bulk_g2s(w_s(s), weights + a0, cnt * static_cast<uint32_t>(sizeof(WT)), &full_bar[s],
               pol_stream);
    if (off_bytes) bulk_g2s(off_s(s), offsets + r0, off_bytes, &full_bar[s], pol_stream);
  };

  if (tid == 0) {
    for (int64_t j = 0; j < STAGES - 1 && j < my_n; ++j) issue(j);
  }

  acc2 acc;

  for (int64_t j = 0; j < my_n; ++j) {
    const tile_t t   = P.tiles[blockIdx.x + j * stride];
    const int s      = static_cast<int>(j % STAGES);
    const uint32_t ph = static_cast<uint32_t>((j / STAGES) & 1);
    const int lead   = static_cast<int>(t.edge_first & 3);
    const int cnt    = (lead + t.num_edges + 3) & ~3;
    const int pb     = static_cast<int>(j & 1);

    mbar_wait(&full_bar[s], ph);

    const int32_t* is = idx_s(s);
    const WT* ws      = w_s(s);
    RT* ps = S::inplace ? reinterpret_cast<RT*>(idx_s(s)) : prod_s(pb);

    if (t.kind == 0) {
      // ---- gather phase: 4 edges per thread per step, products to shared memory ----
      for (int i = tid * 4; i < cnt; i += THREADS * 4) {
        int4 id = *reinterpret_cast<const int4*>(is + i);
        RT v0 = ld_gather(x + id.x), v1 = ld_gather(x + id.y);
        RT v2 = ld_gather(x + id.z), v3 = ld_gather(x + id.w);
        if (WEIGHTED) {
          v0 *= static_cast<RT>(ws[i]);
          v1 *= static_cast<RT>(ws[i + 1]);
          v2 *= static_cast<RT>(ws[i + 2]);
          v3 *= static_cast<RT>(ws[i + 3]);
        }
        ps[i] = v0; ps[i + 1] = v1; ps[i + 2] = v2; ps[i + 3] = v3;
      }
    } else {
      // ---- chunk of a long row: block-wide reduction, no shared-memory round trip ----
      double part = 0.0;
      const int hi = lead + t.num_edges;
      for (int i = tid * 4; i < cnt; i += THREADS * 4) {
        int4 id = *reinterpret_cast<const int4*>(is + i);
        RT v0 = ld_gather(x + id.x), v1 = ld_gather(x + id.y);
        RT v2 = ld_gather(x + id.z), v3 = ld_gather(x + id.w);
        if (WEIGHTED) {
          v0 *= static_cast<RT>(ws[i]);
          v1 *= static_cast<RT>(ws[i + 1]);
          v2 *= static_cast<RT>(ws[i + 2]);
          v3 *= static_cast<RT>(ws[i + 3]);
        }
        RT sum4 = RT(0);
        if (i >= lead && i < hi) sum4 += v0;
        if (i + 1 >= lead && i + 1 < hi) sum4 += v1;
        if (i + 2 >= lead && i + 2 < hi) sum4 += v2;
        if (i + 3 >= lead && i + 3 < hi) sum4 += v3;
        part += static_cast<double>(sum4);
      }
      part = warp_sum(part);
      if (lane == 0) warp_part[pb][warp] = part;
    }

    __syncthreads();
    // every thread has finished tile j-1 entirely and the gather reads of tile j:
    // ring slot (j-1) % STAGES is free again
    if (tid == 0 && j + STAGES - 1 < my_n) issue(j + STAGES - 1);

    if (t.kind == 0) {
      // ---- row reduction + fused epilogue ----
      const bool staged = t.num_rows <= kRowStageCap;
      const ET* offp;
      if (staged) {
        constexpr int per16 = 16 / sizeof(ET);
        offp                = off_s(s) + (t.row_first & (per16 - 1));
      } else {
        offp = offsets + t.row_first;
      }
      const int64_t a0 = t.edge_first - lead;
      const int nrows  = t.num_rows;
      const int avg    = t.num_edges / (nrows > 0 ? nrows : 1);

      auto reduce_rows = [&](auto G_) {
        constexpr int G      = decltype(G_)::value;
        constexpr int groups = THREADS / G;
        const int gid = tid / G, l = tid % G;
        for (int r = gid; r < nrows; r += groups) {
          const int b = static_cast<int>(static_cast<int64_t>(offp[r]) - a0);
          const int e = static_cast<int>(static_cast<int64_t>(offp[r + 1]) - a0);
          RT sum      = RT(0);
          for (int i = b + l; i < e; i += G) sum += ps[i];
#pragma unroll
          for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
          if (l == 0) epi(static_cast<int64_t>(t.row_first) + r, sum, acc);
        }
      };
      // uniform per tile; rounds of a group loop are padded to the warp so that the shuffles
      // always run with a full mask
      auto reduce_rows_padded = [&](auto G_) {
        constexpr int G      = decltype(G_)::value;
        constexpr int groups = THREADS / G;
        constexpr int gpw    = 32 / G;  // groups per warp
        const int gid = tid / G, l = tid % G;
        const int rounds = (nrows + groups - 1) / groups;
        for (int k = 0; k < rounds; ++k) {
          const int r    = k * groups + gid;
          const bool act = r < nrows;
          int b = 0, e = 0;
          if (act) {
            b = static_cast<int>(static_cast<int64_t>(offp[r]) - a0);
            e = static_cast<int>(static_cast<int64_t>(offp[r + 1]) - a0);
          }
          RT sum = RT(0);
          for (int i = b + l; i < e; i += G) sum += ps[i];
#pragma unroll
          for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
          if (act && l == 0) epi(static_cast<int64_t>(t.row_first) + r, sum, acc);
        }
        (void)gpw;
      };
      (void)reduce_rows;
      if (avg < 3) reduce_rows_padded(std::integral_constant<int, 1>{});
      else if (avg < 6) reduce_rows_padded(std::integral_constant<int, 2>{});
      else if (avg < 12) reduce_rows_padded(std::integral_constant<int, 4>{});
      else if (avg < 24) reduce_rows_padded(std::integral_constant<int, 8>{});
      else if (avg < 48) reduce_rows_padded(std::integral_constant<int, 16>{});
      else reduce_rows_padded(std::integral_constant<int, 32>{});
    } else {
      if (warp == 0) {
        double v = (lane < THREADS / 32) ? warp_part[pb][lane] : 0.0;
        v        = warp_sum(v);
        if (lane == 0) P.chunk_partials[t.aux] = v;
      }
    }
  }

  // rows without edges: epilogue only (per_v...cuh:553-559 fills them with init)
  for (int64_t r = P.zero_first + static_cast<int64_t>(blockIdx.x) * THREADS + tid;
       r < P.num_rows; r += static_cast<int64_t>(gridDim.x) * THREADS) {
    epi(r, RT(0), acc);
  }

  if (P.cta_partials) {
    double a = block_sum(acc.a, red_scratch);
    double b = block_sum(acc.b, red_scratch);
    if (tid == 0) {
      P.cta_partials[blockIdx.x]           = a;
      P.cta_partials[P.slots + blockIdx.x] = b;
    }
  }
}

// ------------------------------------------------------------------------------------------
// tail kernel: combine the chunk partials of the long rows (deterministic order), run their
// epilogue, then the last block to finish folds all per-CTA partial sums into the iteration
// state and applies the driver's stop test.
// ------------------------------------------------------------------------------------------
struct tail_params {
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
  const int64_t warps_total = static_cast<int64_t>(gridDim.x) * (THREADS / 32);
  for (int64_t r = static_cast<int64_t>(blockIdx.x) * (THREADS / 32) + warp; r < P.num_long_rows;
       r += warps_total) {
    const int64_t c0 = P.long_row_chunk_first[r], c1 = P.long_row_chunk_first[r + 1];
    double s = 0.0;
    for (int64_t c = c0 + lane; c < c1; c += 32) s += P.chunk_partials[c];
    s = warp_sum(s);
    if (lane == 0) epi(r, static_cast<RT>(s), acc);
  }
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
