// Per-row epilogues fused into the pull kernel.  Each one stands for the vertex passes the
// reference runs as separate thrust launches around per_v_transform_reduce_incoming_e.
// Interface (all __device__):
//   setup(state)                 once per thread at kernel start (reads the iteration scalars)
//   kStage / stage_src(k)        per-row input arrays the TMA producer stages with each tile
//   pre_t from_staged(s0,s1,r,row)  the row's inputs from the staged copies (main kernel)
//   pre_t prefetch(row)          the row's inputs straight from global (tail / fallback kernels)
//   finish(row, sum, pre, acc)   write the row's outputs, add to the per-thread partial sums
//   operator()(row, sum, acc)    = finish(row, sum, prefetch(row), acc)
#pragma once

#include <cstdint>

namespace b200 {

// per-thread accumulators of the fused epilogues (two scalars per iteration)
struct acc2 {
  double a{0.0};  // PageRank: dangling mass of the NEW vector; Katz/eigen: sum of squares
  double b{0.0};  // L1 difference |new - old|
};

// device-resident iteration state shared by all kernels of a driver loop
struct iter_state {
  double a;          // finalised `a` of the last completed iteration
  double b;          // finalised `b`
  double aux;        // driver-specific
  int64_t iterations;
  int32_t done;
  int32_t converged;
  uint32_t ticket;   // last-block election of the tail kernel
  uint32_t pad;
};

// raw: y[row] = init + scale * sum   (b200_spmv_pull, multi-GPU partial sums)
template <typename RT>
struct epi_raw {
  RT* y;
  RT scale;
  RT init;
  struct pre_t {};
  static constexpr int kStage = 0;
  __device__ __forceinline__ const RT* stage_src(int) const { return nullptr; }
  __device__ __forceinline__ pre_t from_staged(const RT*, const RT*, int, int64_t) const { return {}; }
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ pre_t prefetch(int64_t) const { return {}; }
  __device__ __forceinline__ void finish(int64_t row, RT sum, pre_t const&, acc2&) const
  {
    y[row] = init + scale * sum;
  }
  // streaming variants (sell.cuh): inputs / outputs that are touched once per iteration bypass
  // the L2 replacement order (ld.cs / st.cs); outputs of rows < hot_rows are gathered by the next
  // iteration and are stored normally
  __device__ __forceinline__ pre_t prefetch_stream(int64_t) const { return {}; }
  __device__ __forceinline__ void finish_stream(int64_t row, RT sum, pre_t const&, acc2&, int64_t) const
  {
    __stcs(y + row, init + scale * sum);
  }
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    finish(row, sum, prefetch(row), acc);
  }
};

// PageRank (pagerank_impl.cuh:209-292 fused).  The gathered vector is
// x[v] = pr[v] * alpha / out_weight[v], so `sum` already is alpha * sum(pr / outw * w).
template <typename RT>
struct epi_pagerank {
  RT* pr;           // in: old, out: new
  RT* x_next;       // next gathered vector
  const RT* scale;  // alpha / out_weight, or -1 for dangling vertices (out_weight == 0)
  const RT* pers;   // personalization value / personalization sum (dense), or nullptr
  RT alpha;
  RT one_minus_alpha;
  RT num_vertices;  // static_cast<RT>(number_of_vertices)
  RT unvarying;     // set by setup()
  RT pers_factor;   // (dangling * alpha + (1 - alpha)) when personalized
  struct pre_t {
    RT old, sc, pv;
  };
  // pr[] and scale[] of a tile's rows are staged into shared memory by the TMA producer
  static constexpr int kStage = 2;
  __device__ __forceinline__ const RT* stage_src(int k) const { return k == 0 ? pr : scale; }
  __device__ __forceinline__ pre_t from_staged(const RT* s0, const RT* s1, int r, int64_t row) const
  {
    pre_t p;
    p.old = s0[r];
    p.sc  = s1[r];
    p.pv  = pers ? pers[row] : RT(0);
    return p;
  }
  __device__ __forceinline__ void setup(const iter_state* st)
  {
    RT dangling = static_cast<RT>(st->a);
    RT base     = dangling * alpha + one_minus_alpha;
    if (pers) {
      unvarying   = RT(0);
      pers_factor = base;
    } else {
      unvarying   = base / num_vertices;
      pers_factor = RT(0);
    }
  }
  __device__ __forceinline__ pre_t prefetch(int64_t row) const
  {
    pre_t p;
    p.old = pr[row];
    p.sc  = scale[row];
    p.pv  = pers ? pers[row] : RT(0);
    return p;
  }
  __device__ __forceinline__ void finish(int64_t row, RT sum, pre_t const& p, acc2& acc) const
  {
    RT v = sum + unvarying;
    if (pers) v += pers_factor * p.pv;
    pr[row] = v;
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(p.old));
    if (p.sc < RT(0)) {
      acc.a += static_cast<double>(v);
      x_next[row] = RT(0);
    } else {
      x_next[row] = v * p.sc;
    }
  }
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    finish(row, sum, prefetch(row), acc);
  }
  __device__ __forceinline__ pre_t prefetch_stream(int64_t row) const
  {
    pre_t p;
    p.old = __ldcs(pr + row);
    p.sc  = __ldcs(scale + row);
    p.pv  = pers ? __ldcs(pers + row) : RT(0);
    return p;
  }
  __device__ __forceinline__ void finish_stream(int64_t row, RT sum, pre_t const& p, acc2& acc,
                                                int64_t hot_rows) const
  {
    RT v = sum + unvarying;
    if (pers) v += pers_factor * p.pv;
    __stcs(pr + row, v);
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(p.old));
    RT xn = RT(0);
    if (p.sc < RT(0)) acc.a += static_cast<double>(v);
    else xn = v * p.sc;
    if (row < hot_rows) x_next[row] = xn;
    else __stcs(x_next + row, xn);
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
  struct pre_t {
    RT old, b;
  };
  static constexpr int kStage = 1;
  __device__ __forceinline__ const RT* stage_src(int) const { return old_c; }
  __device__ __forceinline__ pre_t from_staged(const RT* s0, const RT*, int r, int64_t row) const
  {
    pre_t p;
    p.old = s0[r];
    p.b   = betas ? betas[row] : beta;
    return p;
  }
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ pre_t prefetch(int64_t row) const
  {
    pre_t p;
    p.old = old_c[row];
    p.b   = betas ? betas[row] : beta;
    return p;
  }
  __device__ __forceinline__ void finish(int64_t row, RT sum, pre_t const& p, acc2& acc) const
  {
    RT v       = alpha * sum + p.b;
    new_c[row] = v;
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(p.old));
    acc.a += static_cast<double>(v) * static_cast<double>(v);
  }
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    finish(row, sum, prefetch(row), acc);
  }
  __device__ __forceinline__ pre_t prefetch_stream(int64_t row) const
  {
    pre_t p;
    p.old = __ldcs(old_c + row);
    p.b   = betas ? __ldcs(betas + row) : beta;
    return p;
  }
  __device__ __forceinline__ void finish_stream(int64_t row, RT sum, pre_t const& p, acc2& acc,
                                                int64_t hot_rows) const
  {
    RT v = alpha * sum + p.b;
    if (row < hot_rows) new_c[row] = v;
    else __stcs(new_c + row, v);
    acc.b += fabs(static_cast<double>(v) - static_cast<double>(p.old));
    acc.a += static_cast<double>(v) * static_cast<double>(v);
  }
};

// eigenvector (eigenvector_centrality_impl.cuh:90-135): new = sum; L2 norm accumulated here,
// normalisation + difference in a separate vertex pass.
template <typename RT>
struct epi_eigen {
  RT* new_c;
  struct pre_t {};
  static constexpr int kStage = 0;
  __device__ __forceinline__ const RT* stage_src(int) const { return nullptr; }
  __device__ __forceinline__ pre_t from_staged(const RT*, const RT*, int, int64_t) const { return {}; }
  __device__ __forceinline__ void setup(const iter_state*) {}
  __device__ __forceinline__ pre_t prefetch(int64_t) const { return {}; }
  __device__ __forceinline__ void finish(int64_t row, RT sum, pre_t const&, acc2& acc) const
  {
    new_c[row] = sum;
    acc.a += static_cast<double>(sum) * static_cast<double>(sum);
  }
  __device__ __forceinline__ void operator()(int64_t row, RT sum, acc2& acc) const
  {
    finish(row, sum, prefetch(row), acc);
  }
  __device__ __forceinline__ pre_t prefetch_stream(int64_t) const { return {}; }
  __device__ __forceinline__ void finish_stream(int64_t row, RT sum, pre_t const&, acc2& acc,
                                                int64_t) const
  {
    new_c[row] = sum;   // normalised (read again) by the driver's vertex pass
    acc.a += static_cast<double>(sum) * static_cast<double>(sum);
  }
};

}  // namespace b200
