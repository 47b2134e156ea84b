// Device-side text parser for edge-list files: the per-thread bodies, written once as
// __host__ __device__ code so that the CPU test-suite can run exactly the statements the kernels
// run (ingest_gpu.cu: the kernels, the driver, and a host emulator of the grid).
//
// Work split.  The entries of the file (everything after the Matrix-Market header / the skipped
// CSV rows) are cut into TILES of kTile bytes, one CTA each, and every tile into CHUNKS of kChunk
// bytes, one thread each.  A line belongs to the chunk its first byte lies in; a thread parses
// the lines that start in its chunk, reading past the chunk (and the tile) as far as the line
// goes, at most kLook bytes.  Pass 1 counts the entries (file entries, and stored entries with
// the mirrored copies of a symmetric file); an exclusive scan turns the counts into output
// positions; pass 2 parses again and writes.  The output order is the file order with the
// mirrored entry right after its entry - the order of the reference's mm_to_coo
// (cpp/tests/utilities/matrix_market_file_utilities.cu:166-215) and of the host reader.
//
// Values.  A decimal number with at most 15 significant digits and a decimal exponent within
// +-22 converts exactly (integer mantissa < 2^53, power of ten exact in double, one correctly
// rounded multiplication or division), so it is bit-identical to strtod / fscanf("%lg").
// Anything else (more digits, large exponents, inf / nan) is marked SLOW: the driver re-parses
// those few values on the host with std::from_chars and patches them in.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define B200_HD __host__ __device__ __forceinline__
#else
#define B200_HD inline
#endif

namespace b200 {
namespace ingest {

// 260 = 65 words: the 32 lanes of a warp, each walking its own chunk byte by byte, then touch 32
// different shared-memory banks (256 would put them all on one)
constexpr int kChunk       = 260;                 // bytes of text per thread
constexpr int kTileThreads = 256;
constexpr int kTile        = kChunk * kTileThreads;  // 65 KB of text per CTA
static_assert(kTile % 16 == 0 && (16 + kTile + 1024) % 16 == 0, "tiles are staged with 128-bit loads");
constexpr int kLook        = 1024;                // a line may run this far past its chunk
constexpr int kTileBuf     = 16 + kTile + kLook;  // [16 bytes before the tile][tile][lookahead]

enum : uint32_t {
  kErrSyntax   = 1,   // a line that is neither blank, a comment, nor `id id [value]`
  kErrNegative = 2,   // an id below the first valid one (0 in a 1-based file)
  kErrLong     = 4,   // a line longer than kLook bytes
};

struct gpu_options {
  int32_t mtx;          // 1-based ids, '%' comment lines
  int32_t mirror;       // store (c, r) right after every (r, c) with r != c
  int32_t has_values;   // a third column is expected
  int32_t want_values;  // ... and stored
  int32_t delimiter;    // CSV field separator (a blank means any run of blanks)
  int32_t vertex_bytes; // 4 | 8
  int32_t weight_bytes; // 4 | 8
  int32_t pad;
};

struct parsed_line {
  int64_t r, c;
  double v;
  int32_t kind;   // 0 not an entry (blank / comment), 1 entry, 2 entry with a SLOW value
  uint32_t err;
};

B200_HD bool gp_is_space(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
B200_HD bool gp_is_digit(char c) { return static_cast<unsigned>(c - '0') <= 9u; }

B200_HD const char* gp_skip_spaces(const char* p, const char* e)
{
  while (p < e && gp_is_space(*p)) ++p;
  return p;
}

// unsigned decimal id with an optional '+'; false on anything else (including a '-' sign: the
// host reader reports those as negative ids, the driver maps both to an error)
B200_HD bool gp_parse_id(const char*& p, const char* e, int64_t& out, bool& negative)
{
  p        = gp_skip_spaces(p, e);
  negative = false;
  if (p < e && *p == '+') ++p;
  else if (p < e && *p == '-') { negative = true; ++p; }
  const char* d0 = p;
  uint64_t v     = 0;
  while (p < e && gp_is_digit(*p) && p - d0 < 18) v = v * 10 + static_cast<unsigned>(*p++ - '0');
  if (p == d0 || (p < e && gp_is_digit(*p))) return false;   // no digits, or more than 18
  out = static_cast<int64_t>(v);
  return true;
}

B200_HD double gp_pow10(int k)
{
  // exact powers of ten 10^0 .. 10^22
  const double t[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                        1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  return t[k];
}

// 0: exact value in v;  1: well-formed but outside the exact range (SLOW);  2: syntax error.
// On 0 and 1, p is left after the number.
B200_HD int gp_parse_value(const char*& p, const char* e, double& v)
{
  p = gp_skip_spaces(p, e);
  bool neg = false;
  if (p < e && (*p == '+' || *p == '-')) { neg = *p == '-'; ++p; }
  const char* start = p;
  uint64_t mant     = 0;
  int sig = 0, dropped_int = 0, frac = 0;
  bool any = false, inexact = false;
  while (p < e && gp_is_digit(*p)) {
    any = true;
    if (sig < 19) {
      if (mant != 0 || *p != '0') { mant = mant * 10 + static_cast<unsigned>(*p - '0'); ++sig; }
    } else {
      ++dropped_int;
      inexact = true;
    }
    ++p;
  }
  if (p < e && *p == '.') {
    ++p;
    while (p < e && gp_is_digit(*p)) {
      any = true;
      if (sig < 19) {
        if (mant != 0 || *p != '0') { mant = mant * 10 + static_cast<unsigned>(*p - '0'); ++sig; }
        ++frac;
      } else if (*p != '0') {
        inexact = true;
      }
      ++p;
    }
  }
  if (!any) {
    // inf / nan / anything alphabetic: let the host decide (it rejects what from_chars rejects)
    p = start;
    while (p < e && !gp_is_space(*p) && *p != ',' && *p != ';') ++p;
    return p > start ? 1 : 2;
  }
  int ex = 0;
  if (p < e && (*p == 'e' || *p == 'E')) {
    const char* q = p + 1;
    bool eneg     = false;
    if (q < e && (*q == '+' || *q == '-')) { eneg = *q == '-'; ++q; }
    if (q < e && gp_is_digit(*q)) {
      int x = 0;
      while (q < e && gp_is_digit(*q)) { if (x < 100000) x = x * 10 + (*q - '0'); ++q; }
      ex = eneg ? -x : x;
      p  = q;
    }
    // an 'e' without digits is not part of the number (from_chars stops before it)
  }
  if (mant == 0 && !inexact) { v = neg ? -0.0 : 0.0; return 0; }
  const int e10 = ex - frac + dropped_int;
  if (inexact || mant >= (1ull << 53) || e10 < -22 || e10 > 22) return 1;
  double d = static_cast<double>(mant);
  d        = e10 < 0 ? d / gp_pow10(-e10) : d * gp_pow10(e10);
  v        = neg ? -d : d;
  return 0;
}

// one line [p, e): e is the position of its '\n' (or the end of the text)
B200_HD void gp_parse_line(const char* p, const char* e, gpu_options const& o, parsed_line& out)
{
  out.kind = 0;
  out.err  = 0;
  out.v    = 1.0;
  const char* q = gp_skip_spaces(p, e);
  if (q == e || (o.mtx && *q == '%')) return;
  bool neg_r = false, neg_c = false;
  bool ok    = gp_parse_id(q, e, out.r, neg_r);
  if (ok) {
    q = gp_skip_spaces(q, e);
    if (!gp_is_space(static_cast<char>(o.delimiter)) && q < e && *q == static_cast<char>(o.delimiter)) ++q;
    ok = gp_parse_id(q, e, out.c, neg_c);
  }
  int vs = 0;
  if (ok && o.has_values) {
    q = gp_skip_spaces(q, e);
    if (!gp_is_space(static_cast<char>(o.delimiter)) && q < e && *q == static_cast<char>(o.delimiter)) ++q;
    vs = gp_parse_value(q, e, out.v);
    ok = vs != 2;
  }
  if (!ok) { out.err = kErrSyntax; return; }
  if (o.mtx) { --out.r; --out.c; }
  if (neg_r || neg_c || out.r < 0 || out.c < 0) {
    // "-0" is id 0 in a 0-based file for from_chars; everything else negative is an error
    if (!(out.r == 0 && !o.mtx) && neg_r) { out.err = kErrNegative; return; }
    if (!(out.c == 0 && !o.mtx) && neg_c) { out.err = kErrNegative; return; }
    if (out.r < 0 || out.c < 0) { out.err = kErrNegative; return; }
  }
  out.kind = vs == 1 ? 2 : 1;
}

// The tile buffer `buf` holds text[tile_base - 16 .. tile_base + kTile + kLook) with '\n' wherever
// that range leaves [0, total) (so the text starts after a line break and ends with one).
// Calls f(line_begin, line_end, global_offset_of_line) for every line that starts inside chunk
// `thread` of the tile; returns an error mask (kErrLong).
template <typename F>
B200_HD uint32_t gp_for_each_line(const char* buf, int64_t tile_base, int64_t total, int thread, F&& f)
{
  const char* text = buf + 16;                       // text[0] = global offset tile_base
  const int begin  = thread * kChunk;
  int64_t room     = total - tile_base;              // bytes of real text from the tile start on
  const int end    = static_cast<int>(room < begin + kChunk ? (room < begin ? begin : room) : begin + kChunk);
  const int limit  = static_cast<int>(room < kTile + kLook ? room : kTile + kLook);  // readable bytes
  uint32_t err     = 0;
  for (int i = begin; i < end; ++i) {
    if (text[i - 1] != '\n') continue;               // not a line start
    int j         = i;
    const int cap = i + kLook < limit ? i + kLook : limit;
    while (j < cap && text[j] != '\n') ++j;
    if (j == cap && cap < room && text[j] != '\n') { err |= kErrLong; continue; }
    f(text + i, text + j, tile_base + i);
    i = j;                                           // next candidate is j + 1 (after the '\n')
  }
  return err;
}

struct chunk_counts {
  uint32_t file_entries;
  uint32_t out_entries;
  uint32_t err;
};

B200_HD chunk_counts gp_pass1(const char* buf, int64_t tile_base, int64_t total, int thread,
                              gpu_options const& o)
{
  chunk_counts c{0, 0, 0};
  const uint32_t long_lines =
    gp_for_each_line(buf, tile_base, total, thread, [&](const char* p, const char* e, int64_t) {
      parsed_line L;
      gp_parse_line(p, e, o, L);
      c.err |= L.err;
      if (L.kind) {
        ++c.file_entries;
        c.out_entries += (o.mirror && L.r != L.c) ? 2u : 1u;
      }
    });
  c.err |= long_lines;
  return c;
}

struct slow_entry {
  int64_t text_offset;   // where the line starts (offset from the start of the entries)
  int64_t out_index;     // where its value goes (the mirrored copy, if any, is out_index + 1)
};

struct pass2_out {
  void* src;
  void* dst;
  void* weights;
  slow_entry* slow;        // capacity slow_capacity
  unsigned long long* slow_count;
  long long* max_id;
  int64_t slow_capacity;
};

B200_HD void gp_store_id(void* a, int bytes, int64_t i, int64_t v)
{
  if (bytes == 4) static_cast<int32_t*>(a)[i] = static_cast<int32_t>(v);
  else static_cast<int64_t*>(a)[i] = v;
}
B200_HD void gp_store_value(void* a, int bytes, int64_t i, double v)
{
  if (bytes == 4) static_cast<float*>(a)[i] = static_cast<float>(v);
  else static_cast<double*>(a)[i] = v;
}

// `add_slow(slot&)` reserves a slot in the slow list (atomic on the device); `raise_max(v)`
// folds an id into the running maximum.
template <typename AddSlow, typename RaiseMax>
B200_HD void gp_pass2(const char* buf, int64_t tile_base, int64_t total, int thread,
                      gpu_options const& o, int64_t out_first, pass2_out const& out,
                      AddSlow&& add_slow, RaiseMax&& raise_max)
{
  int64_t pos    = out_first;
  int64_t local_max = -1;
  gp_for_each_line(buf, tile_base, total, thread, [&](const char* p, const char* e, int64_t off) {
    parsed_line L;
    gp_parse_line(p, e, o, L);
    if (!L.kind) return;
    const bool mirrored = o.mirror && L.r != L.c;
    gp_store_id(out.src, o.vertex_bytes, pos, L.r);
    gp_store_id(out.dst, o.vertex_bytes, pos, L.c);
    if (mirrored) {
      gp_store_id(out.src, o.vertex_bytes, pos + 1, L.c);
      gp_store_id(out.dst, o.vertex_bytes, pos + 1, L.r);
    }
    if (o.want_values) {
      gp_store_value(out.weights, o.weight_bytes, pos, L.v);
      if (mirrored) gp_store_value(out.weights, o.weight_bytes, pos + 1, L.v);
      if (L.kind == 2) add_slow(slow_entry{off, pos});
    }
    const int64_t m = L.r > L.c ? L.r : L.c;
    local_max       = m > local_max ? m : local_max;
    pos += mirrored ? 2 : 1;
  });
  if (local_max >= 0) raise_max(local_max);
}

}  // namespace ingest
}  // namespace b200
