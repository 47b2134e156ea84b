// b200_spmv_pull_noalloc: the pull SpMV  y[dst] = init + scale * sum_{src->dst} x[src] * w  for
// callers that may not allocate (code under CUDA-graph capture, tight serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the
// reference.
//
// b200_spmv_pull (centrality.cu) builds a launcher per call: tile schedule, banded layout,
// chunk-partial and per-CTA scratch buffers, all from the stream-ordered pool.  This variant
// allocates NOTHING, on the host or on the device.  It is the long-row part of that kernel's
// scheme (a row cut into pieces that different CTAs reduce, the partial sums parked in scratch
// memory and combined in a fixed order by a second launch) applied to EVERY row, with the cuts at
// fixed edge positions so that no schedule has to be built:
//
//   launch 1  one CTA per window of kWindow consecutive CSC positions: products into shared
//             memory, one warp per row for the rows that touch the window.  A row that lies
//             completely inside the window is finished here; the piece of a row that enters the
//             window from the left goes to scratch[2k] (the window's HEAD partial), the piece of a
//             row that starts inside and leaves to the right goes to scratch[2k + 1] (TAIL).
//   launch 2  one thread per row: rows without edges get `init`; rows that cross a window
//             boundary add TAIL of their first window and HEAD of every further window, in
//             window order (deterministic), and apply init / scale.
//
// The scratch is the file-scope __device__ array below: 2 doubles per window, kMaxWindows
// windows = 4 MiB, which every GPU that loads the library holds until the process exits.  Graphs
// with more than kMaxWindows * kWindow = 2^30 edges are rejected.  Every call shares the array and
// it keeps the last call's contents; nothing in it has to start at zero (every slot that launch 2
// reads was written by launch 1 of the same call), but two calls that overlap on the GPU (two
// streams, two host threads, two handles on one device) race on it: serialise them.
#include <mutex>
#include <type_traits>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kWindow        = 4096;      // CSC positions per CTA
constexpr int kNaThreads     = 256;
constexpr int64_t kMaxWindows = int64_t{1} << 18;   // 2^18 windows * 4096 = 2^30 edges
constexpr int kMaxDevices    = 64;

static __device__ __align__(256) double g_noalloc_scratch[2 * kMaxWindows];   // 4 MiB

// In host code the array's name is a host-side stand-in, not a device address: the kernels get
// the address cudaGetSymbolAddress returns, looked up once per device with that device current.
double* g_scratch_addr[kMaxDevices] = {};
std::mutex g_scratch_mutex;

double* scratch_for_device(int device)
{
  B200_EXPECTS(device >= 0 && device < kMaxDevices, kInvalidInput,
               "b200_spmv_pull_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[device]) {
    int cur = -1;
    CUDA_TRY(cudaGetDevice(&cur));
    B200_EXPECTS(cur == device, kInvalidInput,
                 "b200_spmv_pull_noalloc: the handle's device is not the current device");
    double* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    B200_EXPECTS(p != nullptr, kUnknownError, "cudaGetSymbolAddress returned NULL");
    g_scratch_addr[device] = p;
  }
  return g_scratch_addr[device];
}

// largest r in [0, n] with off[r] <= e  (off is non-decreasing, off[0] = 0, off[n] = E > e)
template <typename ET>
__device__ __forceinline__ int64_t row_of_edge(const ET* off, int64_t n, int64_t e)
{
  int64_t lo = 0, hi = n;   // invariant: off[lo] <= e < off[hi]
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (static_cast<int64_t>(off[mid]) <= e) lo = mid; else hi = mid;
  }
  return lo;
}

struct no_weight {};

template <typename RT, typename WT>
__device__ __forceinline__ double edge_product(const RT* x, const int32_t* idx, const WT* w, int64_t e)
{
  if constexpr (std::is_same<WT, no_weight>::value) {
    return static_cast<double>(x[idx[e]]);
  } else {
    return static_cast<double>(x[idx[e]]) * static_cast<double>(w[e]);
  }
}

template <typename ET, typename RT, typename WT>
__global__ void __launch_bounds__(kNaThreads) noalloc_window_kernel(
  const ET* __restrict__ off, const int32_t* __restrict__ idx, const WT* __restrict__ w,
  const RT* __restrict__ x, int64_t num_rows, int64_t num_edges, double scale, double init,
  RT* __restrict__ y, double* __restrict__ scratch)
{
  __shared__ double prod[kWindow];
  __shared__ int64_t s_rows[2];
  const int64_t k  = blockIdx.x;
  const int64_t e0 = k * kWindow;
  const int64_t e1 = (e0 + kWindow < num_edges) ? e0 + kWindow : num_edges;
  const int n      = static_cast<int>(e1 - e0);
  if (threadIdx.x == 0) s_rows[0] = row_of_edge(off, num_rows, e0);
  if (threadIdx.x == 32) s_rows[1] = row_of_edge(off, num_rows, e1 - 1);
  for (int i = threadIdx.x; i < n; i += kNaThreads) prod[i] = edge_product<RT, WT>(x, idx, w, e0 + i);
  __syncthreads();
  const int64_t r0 = s_rows[0], r1 = s_rows[1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t r = r0 + warp; r <= r1; r += kNaThreads / 32) {
    const int64_t rb = static_cast<int64_t>(off[r]), re = static_cast<int64_t>(off[r + 1]);
    if (rb == re) continue;   // rows without edges belong to launch 2
    const int b = static_cast<int>((rb > e0 ? rb : e0) - e0);
    const int e = static_cast<int>((re < e1 ? re : e1) - e0);
    double sum  = 0.0;
    for (int i = b + lane; i < e; i += 32) sum += prod[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) {
      if (rb < e0)       scratch[2 * k] = sum;       // HEAD: the row entered from the left
      else if (re > e1)  scratch[2 * k + 1] = sum;   // TAIL: starts here, continues to the right
      else               y[r] = static_cast<RT>(init + scale * sum);
    }
  }
}

template <typename ET, typename RT>
__global__ void __launch_bounds__(kNaThreads) noalloc_rows_kernel(
  const ET* __restrict__ off, int64_t num_rows, double scale, double init, RT* __restrict__ y,
  const double* __restrict__ scratch)
{
  for (int64_t r = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; r < num_rows;
       r += int64_t(gridDim.x) * blockDim.x) {
    const int64_t rb = static_cast<int64_t>(off[r]), re = static_cast<int64_t>(off[r + 1]);
    if (rb == re) { y[r] = static_cast<RT>(init); continue; }
    const int64_t kb = rb / kWindow, ke = (re - 1) / kWindow;
    if (kb == ke) continue;   // finished by launch 1
    double sum = scratch[2 * kb + 1];
    for (int64_t k = kb + 1; k <= ke; ++k) sum += scratch[2 * k];
    y[r] = static_cast<RT>(init + scale * sum);
  }
}

template <typename ET, typename RT, typename WT>
void launch(b200_handle* h, const edge_partition& ep, const void* w, const void* x, double scale,
            double init, void* y, double* scratch)
{
  const int64_t nr = ep.num_rows, ne = ep.num_edges;
  const ET* off = reinterpret_cast<const ET*>(ep.offsets.data());
  if (ne > 0) {
    const unsigned windows = static_cast<unsigned>((ne + kWindow - 1) / kWindow);
    noalloc_window_kernel<ET, RT, WT><<<windows, kNaThreads, 0, h->stream>>>(
      off, ep.indices.data(), static_cast<const WT*>(w), static_cast<const RT*>(x), nr, ne, scale,
      init, static_cast<RT*>(y), scratch);
    CUDA_TRY(cudaGetLastError());
  }
  if (nr > 0) {
    int64_t blocks = (nr + kNaThreads - 1) / kNaThreads;
    if (blocks > int64_t(h->sm_count) * 32) blocks = int64_t(h->sm_count) * 32;
    noalloc_rows_kernel<ET, RT><<<static_cast<unsigned>(blocks), kNaThreads, 0, h->stream>>>(
      off, nr, scale, init, static_cast<RT*>(y), scratch);
    CUDA_TRY(cudaGetLastError());
  }
}

template <typename ET, typename RT>
void dispatch_w(b200_handle* h, const b200_graph* g, const edge_partition& ep, const void* x,
                double scale, double init, void* y, double* scratch)
{
  if (!g->weighted)
    launch<ET, RT, no_weight>(h, ep, nullptr, x, scale, init, y, scratch);
  else if (g->weight_type == B200_FLOAT32)
    launch<ET, RT, float>(h, ep, ep.weights.data(), x, scale, init, y, scratch);
  else
    launch<ET, RT, double>(h, ep, ep.weights.data(), x, scale, init, y, scratch);
}

}  // namespace
}  // namespace b200

extern "C" int b200_spmv_pull_noalloc(b200_handle_t* h, const b200_graph_t* g, int value_type,
                                      const void* x, double scale, double init, void* y)
{
  using namespace b200;
  try {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "invalid handle");
    B200_EXPECTS(g && x && y, kInvalidInput, "NULL argument");
    B200_EXPECTS(value_type == B200_FLOAT32 || value_type == B200_FLOAT64, kInvalidInput,
                 "value_type must be B200_FLOAT32 or B200_FLOAT64");
    B200_EXPECTS(!g->multi_gpu && g->parts.size() == 1, kInvalidInput,
                 "b200_spmv_pull_noalloc: single-GPU graphs only");
    const edge_partition& ep = g->parts[0];
    B200_EXPECTS(ep.num_edges <= kMaxWindows * kWindow, kInvalidInput,
                 "b200_spmv_pull_noalloc: the graph has more than 2^30 edges, the limit of the "
                 "static scratch array (2^18 windows of 4096 edges); use b200_spmv_pull");
    double* scratch = scratch_for_device(h->device);
    if (value_type == B200_FLOAT32) {
      if (ep.offsets64) dispatch_w<int64_t, float>(h, g, ep, x, scale, init, y, scratch);
      else              dispatch_w<int32_t, float>(h, g, ep, x, scale, init, y, scratch);
    } else {
      if (ep.offsets64) dispatch_w<int64_t, double>(h, g, ep, x, scale, init, y, scratch);
      else              dispatch_w<int32_t, double>(h, g, ep, x, scale, init, y, scratch);
    }
    return kSuccess;
  } catch (error const& e) {
    set_last_error(e.what());
    return e.code;
  } catch (std::exception const& e) {
    set_last_error(e.what());
    return kUnknownError;
  }
}
