// Shared internals of libcugraph_b200.so: error plumbing, device buffers, handle / graph
// objects.  Nothing here is exported; the C ABI is include/b200_cugraph.h.
#pragma once

#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/b200_cugraph.h"

namespace b200 {

// cugraph_error_code_t values (cpp/include/cugraph_c/error.h:25-33)
enum : int {
  kSuccess = 0,
  kUnknownError,
  kInvalidHandle,
  kAllocError,
  kInvalidInput,
  kNotImplemented,
  kUnsupportedTypeCombination
};

// the reference throws cugraph::logic_error (utilities/error.hpp:33-49); the C boundary maps
// every exception to an error code + message (c_api/utils.hpp:49-53).
struct error : std::runtime_error {
  int code;
  error(int c, std::string const& m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const char* msg);

#define B200_STR2(x) #x
#define B200_STR(x) B200_STR2(x)

#define B200_EXPECTS(cond, code, msg)                 \
  do {                                                \
    if (!(cond)) throw ::b200::error((code), (msg));  \
  } while (0)

#define CUDA_TRY(call)                                                                     \
  do {                                                                                     \
    cudaError_t e__ = (call);                                                              \
    if (e__ != cudaSuccess) {                                                              \
      char buf__[512];                                                                     \
      snprintf(buf__, sizeof buf__, "CUDA error %s at %s:%d: %s", cudaGetErrorName(e__),   \
               __FILE__, __LINE__, cudaGetErrorString(e__));                               \
      throw ::b200::error(e__ == cudaErrorMemoryAllocation ? ::b200::kAllocError           \
                                                           : ::b200::kUnknownError,        \
                          buf__);                                                          \
    }                                                                                      \
  } while (0)

#define NCCL_TRY(call)                                                                     \
  do {                                                                                     \
    ncclResult_t r__ = (call);                                                             \
    if (r__ != ncclSuccess) {                                                              \
      char buf__[512];                                                                     \
      snprintf(buf__, sizeof buf__, "NCCL error at %s:%d: %s", __FILE__, __LINE__,         \
               ncclGetErrorString(r__));                                                   \
      throw ::b200::error(::b200::kUnknownError, buf__);                                   \
    }                                                                                      \
  } while (0)

// Owning device buffer (the role rmm::device_uvector plays in the reference).  Allocation and
// release are stream-ordered (cudaMallocAsync / cudaFreeAsync on the handle's stream) out of
// the device's default memory pool, whose release threshold the handle raises so that freed
// blocks stay cached - the role of the rmm pool resource: a graph build or a driver call then
// costs no driver-level cudaMalloc / cudaFree (measured: 95 -> 675 ms build time drift at
// scale 24 without the pool).  `pad` extra elements are zero-filled so that 16-byte bulk
// copies may run past the logical end.
template <typename T>
struct dbuf {
  T* p{nullptr};
  size_t n{0};
  cudaStream_t st{nullptr};
  dbuf() = default;
  dbuf(size_t count, cudaStream_t s, size_t pad_elems = 0) { alloc(count, s, pad_elems); }
  dbuf(dbuf const&)            = delete;
  dbuf& operator=(dbuf const&) = delete;
  dbuf(dbuf&& o) noexcept : p(o.p), n(o.n), st(o.st) { o.p = nullptr; o.n = 0; }
  dbuf& operator=(dbuf&& o) noexcept
  {
    if (this != &o) {
      release();
      p = o.p; n = o.n; st = o.st; o.p = nullptr; o.n = 0;
    }
    return *this;
  }
  ~dbuf() { release(); }
  void alloc(size_t count, cudaStream_t s, size_t pad_elems = 0)
  {
    release();
    n            = count;
    st           = s;
    size_t bytes = (count + pad_elems) * sizeof(T);
    if (bytes == 0) bytes = sizeof(T);
    bytes = (bytes + 255) & ~size_t{255};
    CUDA_TRY(cudaMallocAsync(reinterpret_cast<void**>(&p), bytes, s));
    if (pad_elems) {
      CUDA_TRY(cudaMemsetAsync(reinterpret_cast<char*>(p) + count * sizeof(T), 0,
                               bytes - count * sizeof(T), s));
    }
  }
  void release()
  {
    if (p) {
      // the stream may be gone when a graph outlives its handle: fall back to the blocking call
      if (cudaFreeAsync(p, st) != cudaSuccess) {
        cudaGetLastError();
        cudaFree(p);
      }
    }
    p = nullptr; n = 0;
  }
  T* data() const { return p; }
  size_t size() const { return n; }
};

inline int dtype_size(int t) { return (t == B200_INT32 || t == B200_FLOAT32) ? 4 : 8; }

}  // namespace b200

// ---- per-tile schedule of the pull kernel (built once per edge partition) ----------------
namespace b200 {

// One tile = one unit of work of a CTA in the streaming SpMV kernel (see spmv.cuh).
//  kind 0 ("rows"):  a run of COMPLETE rows [row_first, row_first + num_rows) whose edges are
//                    the contiguous range [edge_first, edge_first + num_edges)
//  kind 1 ("chunk"): a slice of ONE long row (high-degree segment); the partial sum goes to
//                    chunk_partials[aux] and is combined by the tail kernel.
struct alignas(16) tile_t {
  int64_t edge_first;
  int32_t num_edges;
  int32_t row_first;
  int32_t num_rows;   // kind 0: rows in the tile; kind 1: 1
  int32_t aux;        // kind 1: index into chunk_partials; kind 0: unused
  int32_t kind;
  int32_t pad;
};
static_assert(sizeof(tile_t) == 32, "tile_t layout");

}  // namespace b200

struct b200_handle {
  cudaStream_t stream{nullptr};
  bool own_stream{false};
  int device{0};
  int sm_count{148};
  size_t l2_bytes{0};
  // multi-GPU
  ncclComm_t world{nullptr}, row{nullptr}, col{nullptr};
  cudaStream_t comm_stream{nullptr};   // per-partition reductions overlap the next SpMV
  int rank{0}, size{1};
  int row_size{1}, col_size{1};   // R, C
  int row_rank{0}, col_rank{0};   // rank inside row_comm / col_comm
  std::shared_ptr<void> peer;     // mg_peer.cuh: IPC-mapped exchange heap of the box's ranks
};

namespace b200 {

struct banded_layout;  // banded.cuh: the pre-reduction layout derived from a CSC rectangle
struct sell_layout;    // sell.cuh: the round-2 layout (lane owns a segment)

// One CSC (or CSR) rectangle + its kernel schedule.
struct edge_partition {
  int64_t num_rows{0};          // majors in this partition
  int64_t num_edges{0};
  int64_t major_first{0};       // global id of row 0
  int64_t minor_first{0};       // global id of column 0 of the gathered vector
  int64_t num_minors{0};
  bool offsets64{false};
  dbuf<char> offsets;           // int32 or int64 [num_rows + 1]
  dbuf<int32_t> indices;        // minor offsets (minor - minor_first)
  dbuf<char> weights;           // float or double, or empty
  int64_t seg[5]{0, 0, 0, 0, 0};  // [0, n_high, n_high+n_mid, n_nonzero, num_rows]
  bool has_segments{false};
  // schedule (one per tile capacity; built lazily by spmv.cu)
  int32_t sched_tile_edges{0};
  int32_t sched_row_cap{0};
  dbuf<tile_t> tiles;
  int64_t num_tiles{0};
  int64_t num_chunks{0};          // kind-1 tiles
  dbuf<int64_t> long_row_chunk_first;  // [n_long + 1] prefix of chunk counts of the long rows
  int64_t num_long_rows{0};
  // multi-GPU partitions: rows follow the GLOBAL degree order, so local degrees are not
  // monotone; the schedule is then packed greedily and the long rows are listed by id
  bool generic_schedule{false};
  dbuf<int32_t> long_row_ids;     // [n_long] (generic schedule only; else rows [0, n_long))
  int64_t zero_first{0};          // rows [zero_first, num_rows) have no edges
  // the minor range is the concatenation of R vertex partitions, each in descending degree
  // order: offsets of those pieces inside the range ({0, num_minors} on a single GPU).  The
  // kernel's shared-memory hub cache takes the head of every piece.
  std::vector<int64_t> minor_piece_offsets;
  // banded pre-reduction layouts derived from this rectangle, one per result width
  // ([0] 4-byte, [1] 8-byte values), built on first use by the pull launcher (banded.cu)
  std::shared_ptr<banded_layout> banded[2];
  std::shared_ptr<sell_layout> sell[2];
  // DCSC form of the hypersparse segment (multi-GPU partitions, graph_mg.cu step 8b)
  dbuf<char> dcs_offsets;          // int32 / int64 like `offsets`
  dbuf<int32_t> dcs_nzd_vertices;  // global internal ids
  int64_t dcs_nzd_count{0};
  int64_t hypersparse_first{-1};   // local row index of the segment's first row, -1 if none
};

}  // namespace b200

struct b200_graph {
  int vertex_type{B200_INT32};
  int edge_type{B200_INT32};
  int weight_type{B200_FLOAT32};
  bool weighted{false};
  bool store_transposed{true};
  bool renumbered{false};
  bool multi_gpu{false};
  int64_t num_vertices{0};        // global
  int64_t num_edges{0};           // global
  int64_t local_first{0}, local_last{0};  // this rank's vertex partition
  std::vector<int64_t> vertex_partition_offsets;  // MG: size P+1
  std::vector<std::vector<int64_t>> partition_segment_offsets;  // MG: one per edge partition
  b200::dbuf<char> renumber_map;  // vertex_type [local V]: internal -> external
  std::vector<b200::edge_partition> parts;  // SG: 1; MG: col_comm_size
  // MG: the C local partitions again as ONE rectangle (rows = the concatenated major ranges,
  // absolute edge offsets), which the drivers run as a single launch per iteration; the
  // per-partition copies above are what the views and the parity tests read.
  std::vector<b200::edge_partition> fused;      // 0 or 1 entries
  std::vector<int64_t> fused_row_first;         // size C + 1: rows of partition i inside `fused`
  int64_t fused_slice{0};                        // rows per partition slice of `fused` (padded, equal)
  int64_t segment_offsets[8]{};
  int num_segments{0};
  // cached out-weight sums (double precision master copy, local vertices)
  b200::dbuf<double> out_weight_sums;
  bool have_out_weight_sums{false};
  // sorted (external id, internal id) pairs of the local vertices, built on first use
  b200::dbuf<uint64_t> ext_sorted;
  b200::dbuf<uint32_t> ext_internal;
  // HITS companions (hits.cu), built on first use: the same graph with rows = sources (CSR) and,
  // for a weighted graph, a weight-less copy with rows = destinations; internal ids
  std::shared_ptr<b200_graph> hits_csr, hits_csc;
  bool have_ext_map{false};
  // the COO in external ids is not kept; a graph stored the other way round is rebuilt from a
  // decompressed edge list when a pull algorithm needs it (c_api/graph.hpp:44-79 does the same)
};
