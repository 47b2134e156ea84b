// Edge-list ingest (include/b200_cugraph.h, "Edge-list ingest"): Matrix-Market and delimited
// text files -> COO in page-locked host memory.  Host code only; the device enters when the
// caller copies the arrays and calls b200_graph_create_sg.
//
// Replaces the per-entry fscanf loop of cpp/tests/utilities/matrix_market_file_utilities.cu
// (mm_properties :53-116, mm_to_coo :133-222, read_edgelist_from_matrix_market_file :266-322)
// with: the file mapped read-only, a split at line boundaries over the host threads, a
// from_chars parse into per-thread vectors, and a copy into the final arrays at the prefix
// offsets - so entry order (and the mirrored entry right after each off-diagonal entry of a
// symmetric file, :199-215) is the reference's whatever the thread count.
#include <algorithm>
#include <charconv>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cuda_runtime.h>

#include "ingest.cuh"
#include "ingest_gpu.cuh"

namespace b200 {
namespace ingest {
namespace {

// vertex ids: plain decimal digits, the common case, without the from_chars call
inline bool parse_id(const char*& p, const char* e, int64_t& out)
{
  p = skip_spaces(p, e);
  const char* q = p;
  if (q < e && *q == '+') ++q;
  uint64_t v       = 0;
  const char* d0   = q;
  while (q < e && static_cast<unsigned>(*q - '0') <= 9u && q - d0 < 18) v = v * 10 + static_cast<unsigned>(*q++ - '0');
  if (q == d0 || (q < e && static_cast<unsigned>(*q - '0') <= 9u)) return parse_number(p, e, out);  // sign, overflow, junk
  out = static_cast<int64_t>(v);
  p   = q;
  return true;
}

// what one host thread parsed, already in the output types (an id that does not fit V is caught
// through max_id before the arrays are used)
template <typename V, typename W>
struct chunk_out {
  std::vector<V> src, dst;
  std::vector<W> val;
  int64_t file_entries{0};
  int64_t max_id{-1};
  std::string err;
};

// one field separator of a delimited file: the delimiter itself, surrounded by optional blanks
// (a blank delimiter therefore means "any run of blanks", which is what fscanf's ' ' does)
inline void skip_delimiter(const char*& p, const char* e, char delim)
{
  p = skip_spaces(p, e);
  if (!is_space(delim) && p < e && *p == delim) ++p;
}

template <typename V, typename W>
void parse_chunk(const char* begin, const char* end, int64_t first_line, parse_options const& o,
                 chunk_out<V, W>& out)
{
  int64_t line_no = first_line;
  // room for one entry per 12 bytes of text (grows if the lines are shorter), twice when mirrored
  size_t guess = static_cast<size_t>(end - begin) / 12 * (o.mirror ? 2 : 1) + 16;
  out.src.reserve(guess);
  out.dst.reserve(guess);
  if (o.want_values) out.val.reserve(guess);
  for (const char* p = begin; p < end; ++line_no) {
    const char* e    = line_end(p, end);
    const char* next = e < end ? e + 1 : end;
    if (blank(p, e) || (o.mtx && *skip_spaces(p, e) == '%')) {
      p = next;
      continue;
    }
    int64_t r{}, c{};
    double v     = 1.0;
    const char* q = p;
    bool ok       = parse_id(q, e, r);
    if (ok) {
      skip_delimiter(q, e, o.delimiter);
      ok = parse_id(q, e, c);
    }
    if (ok && o.has_values) {
      skip_delimiter(q, e, o.delimiter);
      // short decimals convert exactly with one multiplication or division (ingest_gpu.cuh);
      // everything else goes through from_chars.  Both give the correctly rounded double.
      const char* q2 = q;
      if (gp_parse_value(q2, e, v) == 0)
        q = q2;
      else
        ok = parse_number(q, e, v);
    }
    if (!ok) {
      out.err = "could not read matrix data (line " + std::to_string(line_no + 1) + ")";
      return;
    }
    if (o.mtx) {
      --r;
      --c;
    }
    if (r < 0 || c < 0) {
      out.err = "negative vertex id (line " + std::to_string(line_no + 1) + ")";
      return;
    }
    out.max_id = std::max(out.max_id, std::max(r, c));
    ++out.file_entries;
    out.src.push_back(static_cast<V>(r));
    out.dst.push_back(static_cast<V>(c));
    if (o.want_values) out.val.push_back(static_cast<W>(v));
    if (o.mirror && r != c) {
      out.src.push_back(static_cast<V>(c));
      out.dst.push_back(static_cast<V>(r));
      if (o.want_values) out.val.push_back(static_cast<W>(o.skew ? -v : v));
    }
    p = next;
  }
}

// the threads' pieces, copied side by side into the output array
template <typename V, typename W, typename T>
void concatenate(std::vector<chunk_out<V, W>> const& parts, std::vector<int64_t> const& offsets,
                 std::vector<T> chunk_out<V, W>::*member, void* dst_array, unsigned threads)
{
  auto work = [&](size_t i) {
    auto const& piece = parts[i].*member;
    if (!piece.empty())
      memcpy(static_cast<T*>(dst_array) + offsets[i], piece.data(), piece.size() * sizeof(T));
  };
  std::vector<std::thread> pool;
  for (unsigned t = 1; t < threads; ++t)
    pool.emplace_back([&, t] {
      for (size_t i = t; i < parts.size(); i += threads) work(i);
    });
  for (size_t i = 0; i < parts.size(); i += threads) work(i);
  for (auto& th : pool) th.join();
}

template <typename V, typename W>
b200_host_edgelist* parse_body_t(const char* begin, const char* end, int64_t first_line,
                                 parse_options const& o, int vertex_type, int weight_type,
                                 int threads_arg);

}  // namespace

b200_host_edgelist* parse_body(const char* begin, const char* end, int64_t first_line,
                               parse_options const& o, int vertex_type, int weight_type,
                               int threads_arg)
{
  const bool v32 = vertex_type == B200_INT32, w32 = weight_type == B200_FLOAT32;
  if (v32 && w32) return parse_body_t<int32_t, float>(begin, end, first_line, o, vertex_type, weight_type, threads_arg);
  if (v32) return parse_body_t<int32_t, double>(begin, end, first_line, o, vertex_type, weight_type, threads_arg);
  if (w32) return parse_body_t<int64_t, float>(begin, end, first_line, o, vertex_type, weight_type, threads_arg);
  return parse_body_t<int64_t, double>(begin, end, first_line, o, vertex_type, weight_type, threads_arg);
}

namespace {

template <typename V, typename W>
b200_host_edgelist* parse_body_t(const char* begin, const char* end, int64_t first_line,
                                 parse_options const& o, int vertex_type, int weight_type,
                                 int threads_arg)
{
  unsigned hw      = std::max(1u, std::thread::hardware_concurrency());
  size_t bytes     = static_cast<size_t>(end - begin);
  // an explicit thread count is honoured (tests use it on small files); the default is one
  // thread per 64 KB of text up to the hardware threads
  unsigned threads = threads_arg > 0 ? std::min(static_cast<unsigned>(threads_arg), 1024u)
                                     : static_cast<unsigned>(std::min<size_t>(hw, bytes / 65536 + 1));

  // chunk starts, moved forward to the next line start
  std::vector<const char*> cut(threads + 1, end);
  cut[0] = begin;
  for (unsigned t = 1; t < threads; ++t) {
    const char* p = begin + bytes * t / threads;
    if (p <= cut[t - 1]) {
      cut[t] = cut[t - 1];
      continue;
    }
    const char* e = line_end(p - 1, end);  // p - 1: a cut that falls on a line start stays there
    cut[t]        = e < end ? e + 1 : end;
  }
  auto t0 = std::chrono::steady_clock::now();
  std::vector<chunk_out<V, W>> parts(threads);
  {
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t)
      pool.emplace_back([&, t] { parse_chunk(cut[t], cut[t + 1], 0, o, parts[t]); });
    parse_chunk(cut[0], cut[1], first_line, o, parts[0]);
    for (auto& th : pool) th.join();
  }
  for (unsigned t = 0; t < threads; ++t) {
    if (parts[t].err.empty()) continue;
    if (t > 0) {
      // recount the absolute line for the message
      int64_t before = first_line + std::count(begin, cut[t], '\n');
      chunk_out<V, W> redo;
      parse_chunk(cut[t], cut[t + 1], before, o, redo);
      throw error(kInvalidInput, redo.err);
    }
    throw error(kInvalidInput, parts[t].err);
  }

  std::vector<int64_t> offsets(threads + 1, 0);
  int64_t file_entries = 0, max_id = -1;
  for (unsigned t = 0; t < threads; ++t) {
    offsets[t + 1] = offsets[t] + static_cast<int64_t>(parts[t].src.size());
    file_entries += parts[t].file_entries;
    max_id = std::max(max_id, parts[t].max_id);
  }
  int64_t n = offsets[threads];
  B200_EXPECTS(vertex_type == B200_INT64 || max_id <= INT32_MAX, kInvalidInput,
               "vertex id does not fit vertex_type INT32");

  auto t1 = std::chrono::steady_clock::now();
  auto* el = new b200_host_edgelist{};
  try {
    el->src = host_alloc(static_cast<size_t>(n) * sizeof(V), el->pinned_src);
    el->dst = host_alloc(static_cast<size_t>(n) * sizeof(V), el->pinned_dst);
    if (o.want_values) el->weights = host_alloc(static_cast<size_t>(n) * sizeof(W), el->pinned_weights);
    concatenate(parts, offsets, &chunk_out<V, W>::src, el->src, threads);
    concatenate(parts, offsets, &chunk_out<V, W>::dst, el->dst, threads);
    if (o.want_values) concatenate(parts, offsets, &chunk_out<V, W>::val, el->weights, threads);
    if (getenv("B200_INGEST_VERBOSE")) {
      auto t2 = std::chrono::steady_clock::now();
      fprintf(stderr, "[ingest] %u threads: parse %.1f ms, allocate + copy %.1f ms, %lld edges\n",
              threads, std::chrono::duration<double, std::milli>(t1 - t0).count(),
              std::chrono::duration<double, std::milli>(t2 - t1).count(), static_cast<long long>(n));
    }
    el->info.num_edges    = n;
    el->info.file_entries = file_entries;
    el->info.num_rows     = max_id + 1;
    el->info.num_cols     = max_id + 1;
    el->info.vertex_type  = vertex_type;
    el->info.weight_type  = o.want_values ? weight_type : -1;
    el->info.pinned = el->pinned_src && el->pinned_dst && (!o.want_values || el->pinned_weights);
  } catch (...) {
    b200_host_edgelist_free(el);
    throw;
  }
  return el;
}

}  // namespace

}  // namespace ingest
}  // namespace b200

extern "C" {

int b200_read_matrix_market(const char* path, int vertex_type, int weight_type, int threads,
                            b200_host_edgelist_t** out)
{
  using namespace b200;
  using namespace b200::ingest;
  return guarded([&] {
    B200_EXPECTS(out != nullptr, kInvalidInput, "out is NULL");
    *out = nullptr;
    check_types(vertex_type, weight_type, true);
    file_bytes file(path);
    mtx_header h = read_mtx_header(file);
    parse_options o;
    o.mtx         = true;
    o.mirror      = h.mirror;
    o.has_values  = !h.pattern;
    o.want_values = weight_type >= 0;
    auto* el = parse_body(h.body, file.end(), h.body_line, o, vertex_type,
                          weight_type < 0 ? B200_FLOAT32 : weight_type, threads);
    try {
      check_mtx_counts(h, el->info.file_entries, el->info.num_rows - 1);
    } catch (...) {
      b200_host_edgelist_free(el);
      throw;
    }
    el->info.num_rows     = h.rows;
    el->info.num_cols     = h.cols;
    el->info.is_symmetric = h.symmetric ? 1 : 0;
    el->info.is_pattern   = h.pattern ? 1 : 0;
    *out                  = el;
  });
}

int b200_read_csv_edgelist(const char* path, char delimiter, int skip_rows, int vertex_type,
                           int weight_type, int threads, b200_host_edgelist_t** out)
{
  using namespace b200;
  using namespace b200::ingest;
  return guarded([&] {
    B200_EXPECTS(out != nullptr, kInvalidInput, "out is NULL");
    *out = nullptr;
    check_types(vertex_type, weight_type, true);
    B200_EXPECTS(skip_rows >= 0, kInvalidInput, "skip_rows must be >= 0");
    check_delimiter(delimiter);
    file_bytes file(path);
    int64_t line_no = 0;
    const char* p   = csv_body(file, skip_rows, line_no);
    parse_options o;
    o.delimiter   = delimiter;
    o.has_values  = weight_type >= 0;
    o.want_values = weight_type >= 0;
    *out = parse_body(p, file.end(), line_no, o, vertex_type,
                      weight_type < 0 ? B200_FLOAT32 : weight_type, threads);
  });
}

int b200_host_edgelist_info(const b200_host_edgelist_t* e, b200_host_edgelist_info_t* out)
{
  if (e == nullptr || out == nullptr) {
    b200::set_last_error("edge list or out is NULL");
    return b200::kInvalidInput;
  }
  *out = e->info;
  return b200::kSuccess;
}

const void* b200_host_edgelist_src(const b200_host_edgelist_t* e) { return e ? e->src : nullptr; }
const void* b200_host_edgelist_dst(const b200_host_edgelist_t* e) { return e ? e->dst : nullptr; }
const void* b200_host_edgelist_weights(const b200_host_edgelist_t* e)
{
  return e ? e->weights : nullptr;
}

void b200_host_edgelist_free(b200_host_edgelist_t* e)
{
  if (e == nullptr) return;
  void* arrays[3] = {e->src, e->dst, e->weights};
  bool pinned[3]  = {e->pinned_src, e->pinned_dst, e->pinned_weights};
  for (int i = 0; i < 3; ++i) {
    if (arrays[i] == nullptr) continue;
    if (pinned[i]) (void)cudaHostUnregister(arrays[i]);
    free(arrays[i]);
  }
  delete e;
}

}  // extern "C"
