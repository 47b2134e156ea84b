// Device-side edge-list parser (include/b200_cugraph.h "Edge-list ingest", *_device entry
// points): the file's text is copied to HBM once and parsed there - one CTA per 65 KB tile staged
// through shared memory with 128-bit loads, one thread per 260-byte chunk, two passes (count,
// scan, write).  Per-thread bodies: ingest_gpu.cuh.  HBM-bound byte work: per byte of text the
// kernels read it twice and write ~0.5 bytes of COO.
//
// Anything the kernels flag - a malformed line, a line longer than 1 KB, more out-of-range
// values than the patch list holds - sends the file to the host reader (ingest.cu), which either
// parses it or raises the error with its line number; the result is then uploaded.  So the two
// readers accept the same files and give the same arrays.
//
// b200_debug_ingest_emulate runs the same per-thread bodies over the same tiles on the host (no
// device), for the CPU test-suite.

#include "ingest.cuh"
#include "ingest_gpu.cuh"
#include "prims.cuh"

struct b200_device_edgelist {
  b200_host_edgelist_info_t info{};
  b200::dbuf<char> src, dst, weights;
};

namespace b200 {
namespace ingest {
namespace {

constexpr int64_t kSlowCapacity = 1 << 20;

__device__ __forceinline__ void stage_tile(const char* dtext, int64_t tile_base, char* buf)
{
  // dtext[0..16) is the padding before the text, so buf[k] = dtext[tile_base + k]
  const uint4* g = reinterpret_cast<const uint4*>(dtext + tile_base);
  uint4* s       = reinterpret_cast<uint4*>(buf);
  for (int k = threadIdx.x; k < kTileBuf / 16; k += kTileThreads) s[k] = g[k];
  __syncthreads();
}

__global__ void __launch_bounds__(kTileThreads) ingest_pass1_kernel(
  const char* __restrict__ dtext, int64_t total, const __grid_constant__ gpu_options o,
  uint32_t* __restrict__ counts, uint32_t* __restrict__ tile_file, uint32_t* __restrict__ tile_out,
  uint32_t* __restrict__ err)
{
  extern __shared__ uint4 tile_smem[];
  char* buf               = reinterpret_cast<char*>(tile_smem);
  const int64_t tile_base = static_cast<int64_t>(blockIdx.x) * kTile;
  stage_tile(dtext, tile_base, buf);
  const chunk_counts c = gp_pass1(buf, tile_base, total, threadIdx.x, o);
  counts[static_cast<size_t>(blockIdx.x) * kTileThreads + threadIdx.x] =
    (c.file_entries << 16) | c.out_entries;
  __shared__ uint32_t red_scratch[32];
  const uint32_t f = block_sum<uint32_t>(c.file_entries, red_scratch);
  const uint32_t n = block_sum<uint32_t>(c.out_entries, red_scratch);
  if (threadIdx.x == 0) {
    tile_file[blockIdx.x] = f;
    tile_out[blockIdx.x]  = n;
  }
  if (c.err) atomicOr(err, c.err);
}

__global__ void __launch_bounds__(kTileThreads) ingest_pass2_kernel(
  const char* __restrict__ dtext, int64_t total, const __grid_constant__ gpu_options o,
  const uint32_t* __restrict__ counts, const int64_t* __restrict__ tile_first,
  const __grid_constant__ pass2_out out)
{
  extern __shared__ uint4 tile_smem[];
  char* buf               = reinterpret_cast<char*>(tile_smem);
  const int64_t tile_base = static_cast<int64_t>(blockIdx.x) * kTile;
  stage_tile(dtext, tile_base, buf);
  const uint32_t mine =
    counts[static_cast<size_t>(blockIdx.x) * kTileThreads + threadIdx.x] & 0xffffu;
  // exclusive prefix of `mine` over the block: warp scans by shuffle, warp totals through smem
  __shared__ uint32_t warp_tot[kTileThreads / 32];
  uint32_t inc = mine;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[wid] = inc;
  __syncthreads();
  uint32_t before = inc - mine;
  for (int w = 0; w < wid; ++w) before += warp_tot[w];
  if (mine == 0) return;
  gp_pass2(
    buf, tile_base, total, threadIdx.x, o, tile_first[blockIdx.x] + before, out,
    [&](slow_entry const& s) {
      const unsigned long long k = atomicAdd(out.slow_count, 1ull);
      if (k < static_cast<unsigned long long>(out.slow_capacity)) out.slow[k] = s;
    },
    [&](int64_t m) { atomicMax(out.max_id, static_cast<long long>(m)); });
}

template <typename W>
__global__ void patch_values_kernel(W* weights, const int64_t* idx, const double* val,
                                    const uint8_t* mirrored, int64_t n)
{
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  weights[idx[i]] = static_cast<W>(val[i]);
  if (mirrored[i]) weights[idx[i] + 1] = static_cast<W>(val[i]);
}

gpu_options to_gpu_options(parse_options const& o, int vertex_type, int weight_type)
{
  gpu_options g{};
  g.mtx          = o.mtx;
  g.mirror       = o.mirror;
  g.has_values   = o.has_values;
  g.want_values  = o.want_values;
  g.delimiter    = o.delimiter;
  g.vertex_bytes = vertex_type == B200_INT32 ? 4 : 8;
  g.weight_bytes = weight_type == B200_FLOAT32 ? 4 : 8;
  return g;
}

// re-parse the value of one flagged line on the host (std::from_chars, as the host reader does)
bool host_value(const char* p, const char* e, parse_options const& o, double& v, bool& mirrored)
{
  gpu_options g = to_gpu_options(o, B200_INT64, B200_FLOAT64);
  const char* q = gp_skip_spaces(p, e);
  int64_t r{}, c{};
  bool nr{}, nc{};
  if (!gp_parse_id(q, e, r, nr)) return false;
  q = gp_skip_spaces(q, e);
  if (!is_space(o.delimiter) && q < e && *q == o.delimiter) ++q;
  if (!gp_parse_id(q, e, c, nc)) return false;
  q = gp_skip_spaces(q, e);
  if (!is_space(o.delimiter) && q < e && *q == o.delimiter) ++q;
  mirrored = g.mirror && r != c;
  return parse_number(q, e, v);
}

struct patch_list {
  std::vector<int64_t> idx;
  std::vector<double> val;
  std::vector<uint8_t> mirrored;
};

// false: a flagged line does not hold a number after all (the host reader will word the error)
bool build_patches(const char* body, const char* end, parse_options const& o,
                   std::vector<slow_entry> const& slow, patch_list& out)
{
  for (auto const& s : slow) {
    const char* p = body + s.text_offset;
    const char* e = line_end(p, end);
    double v{};
    bool m{};
    if (!host_value(p, e, o, v, m)) return false;
    out.idx.push_back(s.out_index);
    out.val.push_back(v);
    out.mirrored.push_back(m ? 1 : 0);
  }
  return true;
}

void fill_info(b200_host_edgelist_info_t& info, int64_t n, int64_t file_entries, int64_t max_id,
               parse_options const& o, int vertex_type, int weight_type)
{
  info.num_edges    = n;
  info.file_entries = file_entries;
  info.num_rows     = max_id + 1;
  info.num_cols     = max_id + 1;
  info.vertex_type  = vertex_type;
  info.weight_type  = o.want_values ? weight_type : -1;
  info.pinned       = 0;
  info.parser       = 1;
}

// the host reader's result, uploaded (the path of every file the kernels do not take)
b200_device_edgelist* upload_host_parse(b200_handle* h, const char* body, const char* end,
                                        int64_t first_line, parse_options const& o,
                                        int vertex_type, int weight_type)
{
  b200_host_edgelist* el = parse_body(body, end, first_line, o, vertex_type, weight_type, 0);
  auto* out              = new b200_device_edgelist{};
  try {
    cudaStream_t s = h->stream;
    out->info      = el->info;
    out->info.pinned = 0;
    size_t n       = static_cast<size_t>(el->info.num_edges);
    size_t vb      = n * (vertex_type == B200_INT32 ? 4 : 8);
    size_t wb      = n * (weight_type == B200_FLOAT32 ? 4 : 8);
    out->src.alloc(vb, s);
    out->dst.alloc(vb, s);
    CUDA_TRY(cudaMemcpyAsync(out->src.data(), el->src, vb, cudaMemcpyHostToDevice, s));
    CUDA_TRY(cudaMemcpyAsync(out->dst.data(), el->dst, vb, cudaMemcpyHostToDevice, s));
    if (el->weights) {
      out->weights.alloc(wb, s);
      CUDA_TRY(cudaMemcpyAsync(out->weights.data(), el->weights, wb, cudaMemcpyHostToDevice, s));
    }
    CUDA_TRY(cudaStreamSynchronize(s));
  } catch (...) {
    b200_host_edgelist_free(el);
    delete out;
    throw;
  }
  b200_host_edgelist_free(el);
  return out;
}

b200_device_edgelist* parse_on_device(b200_handle* h, const char* body, const char* end,
                                      int64_t first_line, parse_options const& o, int vertex_type,
                                      int weight_type)
{
  cudaStream_t s      = h->stream;
  const int64_t total = end - body;
  const int64_t tiles = std::max<int64_t>(1, div_up(total, kTile));
  B200_EXPECTS(tiles < (int64_t{1} << 31), kInvalidInput, "file too large for the device parser");
  const gpu_options g = to_gpu_options(o, vertex_type, weight_type);
  const bool verbose  = getenv("B200_INGEST_VERBOSE") != nullptr;
  cudaEvent_t ev[4];
  for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
  struct ev_guard {
    cudaEvent_t* e;
    ~ev_guard() { for (int i = 0; i < 4; ++i) cudaEventDestroy(e[i]); }
  } guard{ev};

  // [16 bytes '\n'][text][ '\n' up to the end of the last tile's buffer ]
  const size_t text_bytes = 16 + static_cast<size_t>(tiles) * kTile + kLook + 16;
  dbuf<char> dtext(text_bytes, s);
  CUDA_TRY(cudaEventRecord(ev[0], s));
  CUDA_TRY(cudaMemsetAsync(dtext.data(), '\n', 16, s));
  CUDA_TRY(cudaMemsetAsync(dtext.data() + 16 + total, '\n', text_bytes - 16 - total, s));
  if (total) CUDA_TRY(cudaMemcpyAsync(dtext.data() + 16, body, total, cudaMemcpyHostToDevice, s));
  CUDA_TRY(cudaEventRecord(ev[1], s));

  dbuf<uint32_t> counts(static_cast<size_t>(tiles) * kTileThreads, s);
  dbuf<uint32_t> tile_file(tiles, s), tile_out(tiles, s);
  dbuf<int64_t> file_first(tiles + 1, s), out_first(tiles + 1, s);
  dbuf<uint32_t> err(1, s);
  CUDA_TRY(cudaMemsetAsync(err.data(), 0, 4, s));
  CUDA_TRY(cudaFuncSetAttribute(ingest_pass1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTileBuf));
  CUDA_TRY(cudaFuncSetAttribute(ingest_pass2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTileBuf));
  ingest_pass1_kernel<<<static_cast<unsigned>(tiles), kTileThreads, kTileBuf, s>>>(
    dtext.data(), total, g, counts.data(), tile_file.data(), tile_out.data(), err.data());
  CUDA_TRY(cudaGetLastError());
  exclusive_scan<uint32_t, int64_t>(tile_file.data(), file_first.data(), tiles, s, true);
  exclusive_scan<uint32_t, int64_t>(tile_out.data(), out_first.data(), tiles, s, true);
  uint32_t h_err = 0;
  int64_t n = 0, file_entries = 0;
  CUDA_TRY(cudaMemcpyAsync(&h_err, err.data(), 4, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&n, out_first.data() + tiles, 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaMemcpyAsync(&file_entries, file_first.data() + tiles, 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaEventRecord(ev[2], s));
  CUDA_TRY(cudaStreamSynchronize(s));
  if (h_err) return upload_host_parse(h, body, end, first_line, o, vertex_type, weight_type);

  auto* out = new b200_device_edgelist{};
  try {
    const size_t vb = static_cast<size_t>(n) * g.vertex_bytes;
    out->src.alloc(vb, s);
    out->dst.alloc(vb, s);
    if (o.want_values) out->weights.alloc(static_cast<size_t>(n) * g.weight_bytes, s);
    dbuf<slow_entry> slow(kSlowCapacity, s);
    dbuf<unsigned long long> slow_count(1, s);
    dbuf<long long> max_id(1, s);
    CUDA_TRY(cudaMemsetAsync(slow_count.data(), 0, 8, s));
    CUDA_TRY(cudaMemsetAsync(max_id.data(), 0xff, 8, s));  // -1
    pass2_out po{out->src.data(), out->dst.data(), out->weights.data(), slow.data(),
                 slow_count.data(), max_id.data(), kSlowCapacity};
    ingest_pass2_kernel<<<static_cast<unsigned>(tiles), kTileThreads, kTileBuf, s>>>(
      dtext.data(), total, g, counts.data(), out_first.data(), po);
    CUDA_TRY(cudaGetLastError());
    unsigned long long h_slow = 0;
    long long h_max           = -1;
    CUDA_TRY(cudaMemcpyAsync(&h_slow, slow_count.data(), 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaMemcpyAsync(&h_max, max_id.data(), 8, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaEventRecord(ev[3], s));
    CUDA_TRY(cudaStreamSynchronize(s));
    bool redo = h_slow > static_cast<unsigned long long>(kSlowCapacity);
    if (!redo && h_slow) {
      std::vector<slow_entry> hs(h_slow);
      CUDA_TRY(cudaMemcpyAsync(hs.data(), slow.data(), h_slow * sizeof(slow_entry),
                               cudaMemcpyDeviceToHost, s));
      CUDA_TRY(cudaStreamSynchronize(s));
      patch_list pl;
      redo = !build_patches(body, end, o, hs, pl);
      if (!redo) {
        dbuf<int64_t> d_idx(h_slow, s);
        dbuf<double> d_val(h_slow, s);
        dbuf<uint8_t> d_m(h_slow, s);
        CUDA_TRY(cudaMemcpyAsync(d_idx.data(), pl.idx.data(), h_slow * 8, cudaMemcpyHostToDevice, s));
        CUDA_TRY(cudaMemcpyAsync(d_val.data(), pl.val.data(), h_slow * 8, cudaMemcpyHostToDevice, s));
        CUDA_TRY(cudaMemcpyAsync(d_m.data(), pl.mirrored.data(), h_slow, cudaMemcpyHostToDevice, s));
        const unsigned grid = static_cast<unsigned>(div_up(static_cast<int64_t>(h_slow), 256));
        if (g.weight_bytes == 4)
          patch_values_kernel<float><<<grid, 256, 0, s>>>(
            reinterpret_cast<float*>(out->weights.data()), d_idx.data(), d_val.data(), d_m.data(),
            static_cast<int64_t>(h_slow));
        else
          patch_values_kernel<double><<<grid, 256, 0, s>>>(
            reinterpret_cast<double*>(out->weights.data()), d_idx.data(), d_val.data(), d_m.data(),
            static_cast<int64_t>(h_slow));
        CUDA_TRY(cudaGetLastError());
        CUDA_TRY(cudaStreamSynchronize(s));
      }
    }
    if (redo || (vertex_type == B200_INT32 && h_max > INT32_MAX)) {
      delete out;
      out = nullptr;
      return upload_host_parse(h, body, end, first_line, o, vertex_type, weight_type);
    }
    fill_info(out->info, n, file_entries, h_max, o, vertex_type, weight_type);
    if (verbose) {
      float up = 0, p1 = 0, p2 = 0;
      cudaEventElapsedTime(&up, ev[0], ev[1]);
      cudaEventElapsedTime(&p1, ev[1], ev[2]);
      cudaEventElapsedTime(&p2, ev[2], ev[3]);
      fprintf(stderr,
              "[ingest gpu] %.1f MB: upload %.2f ms, pass 1 + scans %.2f ms, pass 2 %.2f ms "
              "(%.1f GB/s of text over both passes), %lld edges, %llu patched values\n",
              total / 1e6, up, p1, p2, 2.0 * total / 1e6 / std::max(p1 + p2, 1e-3f),
              static_cast<long long>(n), h_slow);
    }
  } catch (...) {
    delete out;
    throw;
  }
  return out;
}

// ---- host emulation of the grid (CPU test-suite) ---------------------------------------------
b200_host_edgelist* emulate(const char* body, const char* end, int64_t first_line,
                            parse_options const& o, int vertex_type, int weight_type,
                            int64_t slow_capacity)
{
  const int64_t total = end - body;
  const int64_t tiles = std::max<int64_t>(1, div_up(total, kTile));
  const gpu_options g = to_gpu_options(o, vertex_type, weight_type);
  std::vector<char> text(16 + static_cast<size_t>(tiles) * kTile + kLook + 16, '\n');
  if (total) memcpy(text.data() + 16, body, static_cast<size_t>(total));
  std::vector<uint32_t> counts(static_cast<size_t>(tiles) * kTileThreads);
  std::vector<int64_t> out_first(tiles + 1, 0);
  int64_t file_entries = 0;
  uint32_t err         = 0;
  for (int64_t t = 0; t < tiles; ++t) {
    uint32_t tile_out = 0;
    for (int th = 0; th < kTileThreads; ++th) {
      chunk_counts c = gp_pass1(text.data() + t * kTile, t * kTile, total, th, g);
      counts[t * kTileThreads + th] = (c.file_entries << 16) | c.out_entries;
      tile_out += c.out_entries;
      file_entries += c.file_entries;
      err |= c.err;
    }
    out_first[t + 1] = out_first[t] + tile_out;
  }
  if (err) return parse_body(body, end, first_line, o, vertex_type, weight_type, 1);
  const int64_t n = out_first[tiles];
  auto* el        = new b200_host_edgelist{};
  try {
    el->src = host_alloc(static_cast<size_t>(n) * g.vertex_bytes, el->pinned_src);
    el->dst = host_alloc(static_cast<size_t>(n) * g.vertex_bytes, el->pinned_dst);
    if (o.want_values)
      el->weights = host_alloc(static_cast<size_t>(n) * g.weight_bytes, el->pinned_weights);
    std::vector<slow_entry> slow;
    unsigned long long slow_count = 0;
    long long max_id              = -1;
    pass2_out po{el->src, el->dst, el->weights, nullptr, &slow_count, &max_id, slow_capacity};
    for (int64_t t = 0; t < tiles; ++t) {
      uint32_t before = 0;
      for (int th = 0; th < kTileThreads; ++th) {
        const uint32_t mine = counts[t * kTileThreads + th] & 0xffffu;
        if (mine)
          gp_pass2(
            text.data() + t * kTile, t * kTile, total, th, g, out_first[t] + before, po,
            [&](slow_entry const& s) {
              if (static_cast<int64_t>(slow_count++) < slow_capacity) slow.push_back(s);
            },
            [&](int64_t m) { max_id = std::max<long long>(max_id, m); });
        before += mine;
      }
    }
    bool redo = static_cast<int64_t>(slow_count) > slow_capacity;
    patch_list pl;
    if (!redo && slow_count) redo = !build_patches(body, end, o, slow, pl);
    if (redo || (vertex_type == B200_INT32 && max_id > INT32_MAX)) {
      b200_host_edgelist_free(el);
      el = nullptr;
      return parse_body(body, end, first_line, o, vertex_type, weight_type, 1);
    }
    for (size_t i = 0; i < pl.idx.size(); ++i) {
      gp_store_value(el->weights, g.weight_bytes, pl.idx[i], pl.val[i]);
      if (pl.mirrored[i]) gp_store_value(el->weights, g.weight_bytes, pl.idx[i] + 1, pl.val[i]);
    }
    fill_info(el->info, n, file_entries, max_id, o, vertex_type, weight_type);
  } catch (...) {
    if (el) b200_host_edgelist_free(el);
    throw;
  }
  return el;
}

struct opened {
  parse_options o;
  mtx_header mh;
  const char* body{nullptr};
  int64_t first_line{0};
  bool mtx{false};
};

opened open_body(file_bytes const& file, bool csv, char delimiter, int skip_rows, int weight_type)
{
  opened r;
  if (csv) {
    B200_EXPECTS(skip_rows >= 0, kInvalidInput, "skip_rows must be >= 0");
    check_delimiter(delimiter);
    r.body          = csv_body(file, skip_rows, r.first_line);
    r.o.delimiter   = delimiter;
    r.o.has_values  = weight_type >= 0;
    r.o.want_values = weight_type >= 0;
  } else {
    r.mtx           = true;
    r.mh            = read_mtx_header(file);
    r.body          = r.mh.body;
    r.first_line    = r.mh.body_line;
    r.o.mtx         = true;
    r.o.mirror      = r.mh.mirror;
    r.o.has_values  = !r.mh.pattern;
    r.o.want_values = weight_type >= 0;
  }
  return r;
}

void finish_info(opened const& op, b200_host_edgelist_info_t& info)
{
  if (!op.mtx) return;
  check_mtx_counts(op.mh, info.file_entries, info.num_rows - 1);
  info.num_rows     = op.mh.rows;
  info.num_cols     = op.mh.cols;
  info.is_symmetric = op.mh.symmetric ? 1 : 0;
  info.is_pattern   = op.mh.pattern ? 1 : 0;
}

int read_device(b200_handle_t* h, const char* path, bool csv, char delimiter, int skip_rows,
                int vertex_type, int weight_type, b200_device_edgelist_t** out)
{
  return guarded([&] {
    B200_EXPECTS(h != nullptr, kInvalidHandle, "handle is NULL");
    B200_EXPECTS(out != nullptr, kInvalidInput, "out is NULL");
    *out = nullptr;
    check_types(vertex_type, weight_type, true);
    file_bytes file(path);
    opened op = open_body(file, csv, delimiter, skip_rows, weight_type);
    CUDA_TRY(cudaSetDevice(h->device));
    auto* el = parse_on_device(h, op.body, file.end(), op.first_line, op.o, vertex_type,
                               weight_type < 0 ? B200_FLOAT32 : weight_type);
    try {
      finish_info(op, el->info);
    } catch (...) {
      delete el;
      throw;
    }
    *out = el;
  });
}

}  // namespace
}  // namespace ingest
}  // namespace b200

extern "C" {

int b200_read_matrix_market_device(b200_handle_t* h, const char* path, int vertex_type,
                                   int weight_type, b200_device_edgelist_t** out)
{
  return b200::ingest::read_device(h, path, false, ' ', 0, vertex_type, weight_type, out);
}

int b200_read_csv_edgelist_device(b200_handle_t* h, const char* path, char delimiter, int skip_rows,
                                  int vertex_type, int weight_type, b200_device_edgelist_t** out)
{
  return b200::ingest::read_device(h, path, true, delimiter, skip_rows, vertex_type, weight_type, out);
}

int b200_device_edgelist_info(const b200_device_edgelist_t* e, b200_host_edgelist_info_t* out)
{
  if (e == nullptr || out == nullptr) {
    b200::set_last_error("edge list or out is NULL");
    return b200::kInvalidInput;
  }
  *out = e->info;
  return b200::kSuccess;
}
const void* b200_device_edgelist_src(const b200_device_edgelist_t* e) { return e ? e->src.data() : nullptr; }
const void* b200_device_edgelist_dst(const b200_device_edgelist_t* e) { return e ? e->dst.data() : nullptr; }
const void* b200_device_edgelist_weights(const b200_device_edgelist_t* e)
{
  return e ? e->weights.data() : nullptr;
}
void b200_device_edgelist_free(b200_device_edgelist_t* e) { delete e; }

// Test hook (not in the public header): the device parser's per-thread bodies run tile by tile
// on the host.  is_csv = 0: Matrix-Market.  slow_capacity < 0: the library's capacity.
int b200_debug_ingest_emulate(const char* path, int is_csv, char delimiter, int skip_rows,
                              int vertex_type, int weight_type, long long slow_capacity,
                              b200_host_edgelist_t** out)
{
  using namespace b200;
  using namespace b200::ingest;
  return guarded([&] {
    B200_EXPECTS(out != nullptr, kInvalidInput, "out is NULL");
    *out = nullptr;
    check_types(vertex_type, weight_type, true);
    file_bytes file(path);
    opened op = open_body(file, is_csv != 0, delimiter, skip_rows, weight_type);
    auto* el  = emulate(op.body, file.end(), op.first_line, op.o, vertex_type,
                        weight_type < 0 ? B200_FLOAT32 : weight_type,
                        slow_capacity < 0 ? kSlowCapacity : slow_capacity);
    try {
      finish_info(op, el->info);
    } catch (...) {
      b200_host_edgelist_free(el);
      throw;
    }
    *out = el;
  });
}

}  // extern "C"
