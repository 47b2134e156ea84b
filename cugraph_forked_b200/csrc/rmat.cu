// R-MAT edge generator: the bit loop of cugraph::generate_rmat_edgelist
// (cpp/src/generators/generate_rmat_edgelist.cu:75-107) over a counter-based random stream.
// The reference draws 2*scale float uniforms per edge from raft::random (not in the tree, stream
// parity unpinned - SURVEY.md 8c); here random number (edge i, bit, k) is the splitmix64
// finaliser of seed + (first_edge + i) * 2 * scale + 2 * bit + k, so any slice of the stream can
// be generated independently on any rank and reproduced bit-exactly on the CPU
// (oracle/graph_oracle.py: rmat_edgelist).
#include "centrality.cuh"

namespace b200 {
namespace {

__device__ __forceinline__ uint64_t splitmix64(uint64_t x)
{
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z          = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z          = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// float32 in [0, 1): top 24 bits * 2^-24
__device__ __forceinline__ float uniform01(uint64_t counter, uint64_t seed)
{
  return static_cast<float>(splitmix64(counter + seed) >> 40) * 5.9604644775390625e-8f;
}

template <typename VT>
__global__ void __launch_bounds__(256) rmat_kernel(int scale, size_t n, double a_plus_b,
                                                   double a_norm, double c_norm, uint64_t seed,
                                                   int clip_and_flip, uint64_t first_edge, VT* src,
                                                   VT* dst)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    uint64_t base = (first_edge + i) * static_cast<uint64_t>(2 * scale);
    VT s = 0, d = 0;
    for (int bit = scale - 1; bit >= 0; --bit) {
      float r0     = uniform01(base + 2 * bit, seed);
      float r1     = uniform01(base + 2 * bit + 1, seed);
      bool src_bit = static_cast<double>(r0) > a_plus_b;
      bool dst_bit = static_cast<double>(r1) > (src_bit ? c_norm : a_norm);
      if (clip_and_flip && s == d && !src_bit && dst_bit) {
        src_bit = true;
        dst_bit = false;
      }
      s += src_bit ? static_cast<VT>(VT(1) << bit) : VT(0);
      d += dst_bit ? static_cast<VT>(VT(1) << bit) : VT(0);
    }
    src[i] = s;
    dst[i] = d;
  }
}

template <typename RT>
__global__ void __launch_bounds__(256) uniform_kernel(RT* out, size_t n, uint64_t seed)
{
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = static_cast<RT>(uniform01(i, seed));
}

}  // namespace

void generate_rmat(b200_handle* h, int scale, size_t num_edges, double a, double b, double c,
                   uint64_t seed, bool clip_and_flip, uint64_t first_edge, int vertex_type,
                   void* src, void* dst)
{
  B200_EXPECTS(scale >= 1 && scale <= (vertex_type == B200_INT32 ? 30 : 62), kInvalidInput,
               "Invalid input argument: scale too large for vertex_t.");
  B200_EXPECTS((a >= 0.0) && (b >= 0.0) && (c >= 0.0) && (a + b + c <= 1.0), kInvalidInput,
               "Invalid input argument: a, b, c should be non-negative and a + b + c should not "
               "be larger than 1.0.");
  if (num_edges == 0) return;
  double a_plus_b = a + b;
  double a_norm   = (a + b) > 0.0 ? a / (a + b) : 0.0;
  double c_norm   = (1.0 - (a + b)) > 0.0 ? c / (1.0 - (a + b)) : 0.0;
  unsigned grid   = static_cast<unsigned>(std::min<size_t>((num_edges + 255) / 256, size_t(h->sm_count) * 32));
  if (vertex_type == B200_INT32)
    rmat_kernel<int32_t><<<grid, 256, 0, h->stream>>>(scale, num_edges, a_plus_b, a_norm, c_norm, seed,
                                                     clip_and_flip, first_edge,
                                                     static_cast<int32_t*>(src),
                                                     static_cast<int32_t*>(dst));
  else
    rmat_kernel<int64_t><<<grid, 256, 0, h->stream>>>(scale, num_edges, a_plus_b, a_norm, c_norm, seed,
                                                     clip_and_flip, first_edge,
                                                     static_cast<int64_t*>(src),
                                                     static_cast<int64_t*>(dst));
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(h->stream));
}

void uniform_fill(b200_handle* h, int value_type, void* out, size_t n, uint64_t seed)
{
  if (n == 0) return;
  unsigned grid = static_cast<unsigned>(std::min<size_t>((n + 255) / 256, size_t(h->sm_count) * 32));
  if (value_type == B200_FLOAT32)
    uniform_kernel<float><<<grid, 256, 0, h->stream>>>(static_cast<float*>(out), n, seed);
  else
    uniform_kernel<double><<<grid, 256, 0, h->stream>>>(static_cast<double*>(out), n, seed);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(h->stream));
}

}  // namespace b200
