// Hand-written device primitives used by the graph build (no Thrust / CUB): block reductions,
// exclusive scan, stable LSD radix sort.  They replace the thrust::sort / exclusive_scan /
// reduce calls of cpp/src/structure/renumber_edgelist_impl.cuh and
// cpp/src/structure/detail/structure_utils.cuh.
#pragma once

#include "common.cuh"

namespace b200 {

inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ------------------------------------------------------------------------------------------
// warp / block reductions
// ------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_sum(T v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T>
__device__ __forceinline__ T warp_max(T v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    T u = __shfl_xor_sync(0xffffffffu, v, o);
    v   = u > v ? u : v;
  }
  return v;
}

// all threads of the block must call; result valid in every thread. scratch: >= 32 elements.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch)
{
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  T r = (lane < nw) ? scratch[lane] : T(0);
  r   = warp_sum(r);
  return r;
}

// ------------------------------------------------------------------------------------------
// exclusive scan (three-phase, recursive on the block sums); in == out allowed
// ------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems   = 8;
constexpr int kScanChunk   = kScanThreads * kScanItems;

template <typename Tin, typename T>
__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const Tin* in, T* block_sums,
                                                                  size_t n)
{
  __shared__ T scratch[32];
  size_t base = static_cast<size_t>(blockIdx.x) * kScanChunk;
  T s         = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    size_t idx = base + static_cast<size_t>(i) * kScanThreads + threadIdx.x;
    if (idx < n) s += static_cast<T>(in[idx]);
  }
  s = block_sum(s, scratch);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = s;
}

template <typename Tin, typename T>
__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const Tin* in, T* out,
                                                                 const T* block_offsets,
                                                                 size_t n_in, size_t n_out)
{
  __shared__ T warp_tot[32];
  size_t base = static_cast<size_t>(blockIdx.x) * kScanChunk +
                static_cast<size_t>(threadIdx.x) * kScanItems;
  T v[kScanItems];
  T tsum = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n_in) ? static_cast<T>(in[base + i]) : T(0);
    tsum += v[i];
  }
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T u = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += u;
  }
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  if (w == 0) {
    T t  = (lane < kScanThreads / 32) ? warp_tot[lane] : T(0);
    T ti = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T u = __shfl_up_sync(0xffffffffu, ti, o);
      if (lane >= o) ti += u;
    }
    if (lane < kScanThreads / 32) warp_tot[lane] = ti - t;  // exclusive warp offsets
  }
  __syncthreads();
  T run = (block_offsets ? block_offsets[blockIdx.x] : T(0)) + warp_tot[w] + (inc - tsum);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n_out) out[base + i] = run;
    run += v[i];
  }
}

// out[i] = sum_{j<i} in[j] for i < n; with write_total also out[n] = sum of all inputs
// (out must then hold n + 1 elements).  in == out allowed.
template <typename Tin, typename T>
void exclusive_scan(const Tin* in, T* out, size_t n, cudaStream_t s, bool write_total = false)
{
  size_t n_out = n + (write_total ? 1 : 0);
  if (n_out == 0) return;
  size_t nblocks = static_cast<size_t>(div_up(static_cast<int64_t>(n_out), kScanChunk));
  if (nblocks == 1) {
    scan_apply_kernel<Tin, T><<<1, kScanThreads, 0, s>>>(in, out, nullptr, n, n_out);
    CUDA_TRY(cudaGetLastError());
  } else {
    dbuf<T> sums(nblocks, s);
    scan_reduce_kernel<Tin, T><<<nblocks, kScanThreads, 0, s>>>(in, sums.data(), n);
    CUDA_TRY(cudaGetLastError());
    exclusive_scan<T, T>(sums.data(), sums.data(), nblocks, s);
    scan_apply_kernel<Tin, T><<<nblocks, kScanThreads, 0, s>>>(in, out, sums.data(), n, n_out);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(s));  // `sums` is freed on return
  }
}

// ------------------------------------------------------------------------------------------
// stable LSD radix sort, 8-bit digits, key = uint32 / uint64, optional uint32 payload
// ------------------------------------------------------------------------------------------
constexpr int kSortThreads = 256;
constexpr int kSortRounds  = 16;
constexpr int kSortChunk   = kSortThreads * kSortRounds;

template <typename K>
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const K* keys, uint32_t* hist,
                                                                 size_t n, int shift,
                                                                 uint32_t nblocks)
{
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  size_t base = static_cast<size_t>(blockIdx.x) * kSortChunk;
#pragma unroll 4
  for (int r = 0; r < kSortRounds; ++r) {
    size_t idx = base + static_cast<size_t>(r) * kSortThreads + threadIdx.x;
    if (idx < n) atomicAdd(&h[(keys[idx] >> shift) & 255], 1u);
  }
  __syncthreads();
  hist[static_cast<size_t>(threadIdx.x) * nblocks + blockIdx.x] = h[threadIdx.x];
}

// One tile of kSortChunk elements per block.  The tile is first sorted by digit inside shared
// memory (stable: rounds in order, warps in order, lanes ranked with match_any), then written
// out so that consecutive threads store consecutive addresses of the same digit's run - the
// direct register -> global scatter of the first version wrote one 8-byte key per 32-byte
// sector and ran at ~12 % of its HBM roofline (profiles/r1b_ncu_pull_tiles_s24_f32.md).
template <typename K, bool HAS_V>
struct scatter_smem {
  static constexpr size_t keys  = sizeof(K) * kSortChunk;
  static constexpr size_t vals  = HAS_V ? sizeof(uint32_t) * kSortChunk : 0;
  static constexpr size_t cnt   = sizeof(uint32_t) * (kSortThreads / 32) * 256;
  static constexpr size_t small = sizeof(uint32_t) * 256 * 3;  // running, lstart, gbase
  static constexpr size_t total = keys + vals + cnt + small;
};

template <typename K, bool HAS_V>
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(
  const K* kin, K* kout, const uint32_t* vin, uint32_t* vout, const uint32_t* bases, size_t n,
  int shift, uint32_t nblocks)
{
  extern __shared__ __align__(16) unsigned char sort_smem[];
  using S = scatter_smem<K, HAS_V>;
  K* skey            = reinterpret_cast<K*>(sort_smem);
  uint32_t* sval     = reinterpret_cast<uint32_t*>(sort_smem + S::keys);
  uint32_t* warp_cnt = reinterpret_cast<uint32_t*>(sort_smem + S::keys + S::vals);  // [8][256]
  uint32_t* running  = warp_cnt + (kSortThreads / 32) * 256;  // per digit: elements so far
  uint32_t* lstart   = running + 256;                         // per digit: start inside the tile
  uint32_t* gbase    = lstart + 256;                          // per digit: start in the output
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const size_t base = static_cast<size_t>(blockIdx.x) * kSortChunk;
  running[tid] = 0;
  gbase[tid]   = bases[static_cast<size_t>(tid) * nblocks + blockIdx.x];

  K key[kSortRounds];
  uint32_t val[kSortRounds];
  uint16_t lpos[kSortRounds];
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    const size_t idx = base + static_cast<size_t>(r) * kSortThreads + tid;
    key[r] = idx < n ? kin[idx] : K(0);
    val[r] = 0;
    if (HAS_V && idx < n) val[r] = vin[idx];
  }
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
#pragma unroll
    for (int i = 0; i < kSortThreads / 32; ++i) warp_cnt[i * 256 + tid] = 0;
    __syncthreads();
    const bool valid    = base + static_cast<size_t>(r) * kSortThreads + tid < n;
    const int digit     = valid ? static_cast<int>((key[r] >> shift) & 255) : 256;
    const unsigned mask = __match_any_sync(0xffffffffu, digit);
    const int rank      = __popc(mask & ((1u << lane) - 1u));
    if (valid && rank == 0) warp_cnt[w * 256 + digit] = __popc(mask);
    __syncthreads();
    {
      uint32_t run = running[tid];
#pragma unroll
      for (int i = 0; i < kSortThreads / 32; ++i) {
        const uint32_t c        = warp_cnt[i * 256 + tid];
        warp_cnt[i * 256 + tid] = run;
        run += c;
      }
      running[tid] = run;
    }
    __syncthreads();
    lpos[r] = valid ? static_cast<uint16_t>(warp_cnt[w * 256 + digit] + rank) : uint16_t(0);
    __syncthreads();
  }
  // running[] is now the tile's histogram: exclusive scan over the 256 digits
  {
    const uint32_t c = running[tid];
    uint32_t inc     = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += u;
    }
    if (lane == 31) warp_cnt[w] = inc;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int i = 0; i < kSortThreads / 32; ++i)
      if (i < w) off += warp_cnt[i];
    lstart[tid] = off + inc - c;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kSortRounds; ++r) {
    if (base + static_cast<size_t>(r) * kSortThreads + tid < n) {
      const uint32_t p = lstart[(key[r] >> shift) & 255] + lpos[r];
      skey[p]          = key[r];
      if (HAS_V) sval[p] = val[r];
    }
  }
  __syncthreads();
  const uint32_t ntile =
    static_cast<uint32_t>(n - base < static_cast<size_t>(kSortChunk) ? n - base : kSortChunk);
  for (uint32_t j = tid; j < ntile; j += kSortThreads) {
    const K k        = skey[j];
    const uint32_t d = static_cast<uint32_t>((k >> shift) & 255);
    const uint32_t g = gbase[d] + (j - lstart[d]);
    kout[g]          = k;
    if (HAS_V) vout[g] = sval[j];
  }
}

// Sorts n keys (n < 2^32) by the bits [bit_begin, bit_end) of every (begin,end) range given,
// least significant range first.  Returns 0 if the result is in (k0, v0), 1 if in (k1, v1).
template <typename K>
int radix_sort(K* k0, K* k1, uint32_t* v0, uint32_t* v1, size_t n,
               std::vector<std::pair<int, int>> const& bit_ranges, cudaStream_t s)
{
  if (n == 0) return 0;
  B200_EXPECTS(n < (size_t{1} << 32), kInvalidInput, "radix_sort: n must be < 2^32");
  uint32_t nblocks = static_cast<uint32_t>(div_up(static_cast<int64_t>(n), kSortChunk));
  dbuf<uint32_t> hist(static_cast<size_t>(nblocks) * 256, s);
  int cur = 0;
  for (auto const& br : bit_ranges) {
    for (int shift = br.first; shift < br.second; shift += 8) {
      K* kin         = cur ? k1 : k0;
      K* kout        = cur ? k0 : k1;
      uint32_t* vin  = cur ? v1 : v0;
      uint32_t* vout = cur ? v0 : v1;
      radix_hist_kernel<K><<<nblocks, kSortThreads, 0, s>>>(kin, hist.data(), n, shift, nblocks);
      CUDA_TRY(cudaGetLastError());
      exclusive_scan<uint32_t, uint32_t>(hist.data(), hist.data(),
                                         static_cast<size_t>(nblocks) * 256, s);
      if (v0) {
        auto kern = radix_scatter_kernel<K, true>;
        constexpr size_t smem = scatter_smem<K, true>::total;
        CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
        kern<<<nblocks, kSortThreads, smem, s>>>(kin, kout, vin, vout, hist.data(), n, shift, nblocks);
      } else {
        auto kern = radix_scatter_kernel<K, false>;
        constexpr size_t smem = scatter_smem<K, false>::total;
        CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      static_cast<int>(smem)));
        kern<<<nblocks, kSortThreads, smem, s>>>(kin, kout, nullptr, nullptr, hist.data(), n, shift,
                                                 nblocks);
      }
      CUDA_TRY(cudaGetLastError());
      cur ^= 1;
    }
  }
  CUDA_TRY(cudaStreamSynchronize(s));
  return cur;
}

inline int bits_needed(uint64_t max_value)
{
  int b = 0;
  while (max_value) { ++b; max_value >>= 1; }
  return b == 0 ? 1 : b;
}

}  // namespace b200
