// Peer-memory exchange of the multi-GPU iteration (one process per GPU, one NVSwitch box).
//
// Round 1 moved the gathered vector and the partial sums with NCCL calls (grouped send/recv for
// the row-communicator allgatherv: 146 GB/s at 67 MB; C x ncclReduce / one reduce-scatter on the
// column communicator): at 8 GPUs those two steps were 42 % (scale 24) and 57 % (scale 26) of an
// iteration (profiles/r2_notes.md).  Here every rank owns one cudaMalloc'ed HEAP that all ranks
// of the box map through CUDA IPC, and the exchange becomes loads over NVLink:
//   * the local slice of the gathered vector is written into the heap; after a flag barrier among
//     the row peers every rank copies its peers' slices straight out of their heaps;
//   * the SpMV writes its C partial-sum slices into the heap; after a flag barrier among the
//     column peers the owner of a slice SUMS the C copies (its own + C - 1 read over NVLink,
//     coalesced 16-byte loads) in one kernel - reduction and transfer are the same loads, in a
//     fixed order (deterministic, unlike a ring).
// Both heap regions are double-buffered, so one barrier per exchange is enough: a buffer is
// rewritten two iterations later, behind the next barrier of the same kind, which a peer only
// reaches after it has finished reading.
// The barrier is a kernel of one warp: lane t stores this rank's epoch into peer t's flag array
// (st.release.sys after a system fence) and spins on peer t's epoch in the local array
// (ld.acquire.sys).  Every rank runs on its own GPU, so the waits cannot starve one another; a
// wait that exceeds ~10 s sets an error flag instead of hanging (a peer died).
// The reference's counterpart: device_allgatherv / device_reduce over NCCL
// (update_edge_src_dst_property.cuh:335-346, per_v_transform_reduce_incoming_outgoing_e.cuh:915-965).
#pragma once

#include <memory>

#include "epilogues.cuh"
#include "mg_comm.cuh"
#include "prims.cuh"

namespace b200 {

constexpr size_t kPeerHeader  = 4096;   // flag arrays live in the first page of every heap
constexpr int kPeerMaxRanks   = 64;

struct peer_heap {
  int world{1}, rank{0};
  void* base{nullptr};
  size_t bytes{0};
  std::vector<void*> peer_base;     // [world]; peer_base[rank] == base
  uint64_t epoch_row{0}, epoch_col{0}, epoch_scal{0};
  int* d_error{nullptr};            // set by a barrier that timed out
  ~peer_heap()
  {
    for (int p = 0; p < static_cast<int>(peer_base.size()); ++p)
      if (p != rank && peer_base[p]) cudaIpcCloseMemHandle(peer_base[p]);
    if (base) cudaFree(base);
    if (d_error) cudaFree(d_error);
    cudaGetLastError();
  }
  // header of a heap: epochs of the row barrier [0, 512), of the column barrier [512, 1024), of
  // the scalar exchange [1024, 1536); the exchanged scalar pairs, two parities, [2048, 4096)
  static size_t flag_offset(bool col, int comm_rank) { return (col ? 64 : 0) * 8 + size_t(comm_rank) * 8; }
  static size_t scal_flag_offset(int world_rank) { return 1024 + size_t(world_rank) * 8; }
  static size_t scal_val_offset(int parity, int world_rank) { return 2048 + size_t(parity) * 1024 + size_t(world_rank) * 16; }
  template <typename T>
  T* at(int p, size_t byte_offset) const
  {
    return reinterpret_cast<T*>(static_cast<char*>(peer_base[p]) + byte_offset);
  }
};

// collective over h->world: every rank ends with a heap of at least `want` payload bytes
inline peer_heap* peer_heap_get(b200_handle* h, size_t want)
{
  std::vector<int64_t> all = allgather_host(h, h->world, h->size, {static_cast<int64_t>(want)});
  int64_t need = 0;
  for (int64_t v : all) need = std::max(need, v);
  auto* cur = static_cast<peer_heap*>(h->peer.get());
  if (cur && cur->bytes >= static_cast<size_t>(need) + kPeerHeader) return cur;
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  h->peer.reset();
  auto heap   = std::make_shared<peer_heap>();
  heap->world = h->size;
  heap->rank  = h->rank;
  heap->bytes = ((static_cast<size_t>(need) + kPeerHeader + (size_t{2} << 20) - 1) >> 21) << 21;
  // Every rank runs every collective below whatever happens to its own CUDA calls: a failure
  // (no IPC between these processes, no peer access) is carried in `ok` and agreed on at the end,
  // so that all ranks fall back to the NCCL exchange together instead of one of them hanging.
  int64_t ok = 1;
  auto soft  = [&](cudaError_t e) { if (e != cudaSuccess) { ok = 0; cudaGetLastError(); } };
  soft(cudaMalloc(&heap->base, heap->bytes));
  if (ok) soft(cudaMemset(heap->base, 0, kPeerHeader));
  soft(cudaMalloc(reinterpret_cast<void**>(&heap->d_error), sizeof(int)));
  if (heap->d_error) soft(cudaMemset(heap->d_error, 0, sizeof(int)));
  soft(cudaDeviceSynchronize());
  cudaIpcMemHandle_t mine{};
  if (ok) soft(cudaIpcGetMemHandle(&mine, heap->base));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  std::vector<int64_t> words(9);
  std::memcpy(words.data(), &mine, 64);
  words[8] = ok;
  std::vector<int64_t> handles = allgather_host(h, h->world, h->size, words);
  for (int p = 0; p < h->size; ++p) ok = ok && handles[9 * p + 8];
  heap->peer_base.assign(h->size, nullptr);
  for (int p = 0; p < h->size && ok; ++p) {
    if (p == h->rank) { heap->peer_base[p] = heap->base; continue; }
    cudaIpcMemHandle_t theirs;
    std::memcpy(&theirs, handles.data() + 9 * p, 64);
    soft(cudaIpcOpenMemHandle(&heap->peer_base[p], theirs, cudaIpcMemLazyEnablePeerAccess));
  }
  // nobody may signal into a heap that is not mapped everywhere yet
  for (int64_t v : allgather_host(h, h->world, h->size, {ok})) ok = ok && v;
  if (!ok) {
    if (h->rank == 0)
      fprintf(stderr, "[b200] peer-memory exchange unavailable (CUDA IPC / peer access failed on a "
                      "rank): the multi-GPU iteration uses the NCCL exchange\n");
    return nullptr;   // `heap` unmaps and frees what it got
  }
  h->peer = heap;
  return heap.get();
}

// slots[t] = address of this rank's epoch inside peer t's flag array; mine[t] = peer t's epoch
// in the local flag array
__global__ void peer_barrier_kernel(uint64_t* const* slots, const uint64_t* mine, int npeers,
                                    uint64_t epoch, int* error)
{
  const int t = threadIdx.x;
  __threadfence_system();
  if (t < npeers) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(slots[t]), "l"(epoch) : "memory");
    const long long t0 = clock64();
    for (;;) {
      uint64_t v;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(mine + t) : "memory");
      if (v >= epoch) break;
      if (clock64() - t0 > 20000000000ll) { *error = 1; break; }   // ~10 s at 2 GHz
    }
  }
  __threadfence_system();
}

// The two scalars of an iteration (PageRank: dangling mass and L1 difference), summed over ALL
// ranks without a collective call: fold this rank's per-CTA partial sums, store the pair into every
// rank's heap, wait for every rank's pair, add the pairs in rank order (every rank gets the same
// bits) and commit them to the iteration state - one launch of one CTA on the main stream
// instead of fold kernel + ncclAllReduce (on a second stream, which the persistent SpMV kernels
// starve of SMs) + commit kernel.
//   val_slots[t] / flag_slots[t]: where this rank's pair / epoch go inside rank t's heap;
//   vals / flags: the local arrays (indexed by world rank) of this parity
__global__ void __launch_bounds__(256) peer_scalars_kernel(
  const double* partials, int n, double* const* val_slots, uint64_t* const* flag_slots,
  const double* vals, const uint64_t* flags, int world, uint64_t epoch, iter_state* st,
  double threshold, int64_t iteration, int check, int which, int* error)
{
  __shared__ double scratch[32];
  __shared__ double mine[2];
  const int tid = threadIdx.x;
  const bool done = st->done != 0;
  double a = 0.0, b = 0.0;
  if (!done) {
    for (int i = tid; i < n; i += blockDim.x) {
      a += partials[i];
      b += partials[n + i];
    }
  }
  a = block_sum(a, scratch);
  b = block_sum(b, scratch);
  if (tid == 0) { mine[0] = a; mine[1] = b; }
  __syncthreads();
  if (tid < world) {
    double* v = val_slots[tid];
    asm volatile("st.volatile.global.f64 [%0], %1;" ::"l"(v), "d"(mine[0]) : "memory");
    asm volatile("st.volatile.global.f64 [%0], %1;" ::"l"(v + 1), "d"(mine[1]) : "memory");
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag_slots[tid]), "l"(epoch) : "memory");
    const long long t0 = clock64();
    for (;;) {
      uint64_t f;
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(f) : "l"(flags + tid) : "memory");
      if (f >= epoch) break;
      if (clock64() - t0 > 20000000000ll) { *error = 1; break; }
    }
  }
  __syncthreads();
  if (tid == 0 && !done) {
    double sa = 0.0, sb = 0.0;
    for (int t = 0; t < world; ++t) {
      double x0, x1;
      asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(x0) : "l"(vals + 2 * t) : "memory");
      asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(x1) : "l"(vals + 2 * t + 1) : "memory");
      sa += x0;
      sb += x1;
    }
    if (which & 1) st->a = sa;
    if (which & 2) st->b = sb;
    if (iteration > 0) st->iterations = iteration;
    if (check && sb < threshold) {
      st->converged = 1;
      st->done      = 1;
    }
  }
}

// out[i] = sum over the C copies of this rank's slice, in column-rank order (fixed => deterministic)
template <typename RT, int C>
__global__ void __launch_bounds__(256) peer_reduce_kernel(const RT* const* src, RT* out, size_t n)
{
  constexpr int V = 16 / sizeof(RT);
  const RT* s[C];
#pragma unroll
  for (int j = 0; j < C; ++j) s[j] = src[j];
  const size_t nv = n / V;
  for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < nv;
       i += size_t(gridDim.x) * blockDim.x) {
    alignas(16) RT acc[V];
#pragma unroll
    for (int k = 0; k < V; ++k) acc[k] = RT(0);
#pragma unroll
    for (int j = 0; j < C; ++j) {
      const int4 q = __ldcv(reinterpret_cast<const int4*>(s[j]) + i);   // never a stale cached line
      const RT* qv = reinterpret_cast<const RT*>(&q);
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k] += qv[k];
    }
    *reinterpret_cast<int4*>(out + i * V) = *reinterpret_cast<const int4*>(acc);
  }
  for (size_t i = nv * V + blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    RT a = RT(0);
#pragma unroll
    for (int j = 0; j < C; ++j) a += __ldcv(s[j] + i);
    out[i] = a;
  }
}

}  // namespace b200
