// Shared pieces of the edge-list readers: the host reader (ingest.cu) and the device parser
// (ingest_gpu.cu).  See include/b200_cugraph.h "Edge-list ingest".
#pragma once
#include <algorithm>
#include <charconv>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cuda_runtime.h>

#include "../../include/b200_cugraph.h"
#include "common.cuh"

struct b200_host_edgelist {
  b200_host_edgelist_info_t info{};
  void* src{nullptr};
  void* dst{nullptr};
  void* weights{nullptr};
  bool pinned_src{false}, pinned_dst{false}, pinned_weights{false};
};

namespace b200 {
namespace ingest {


// the file, mapped read-only: the parser threads read the page cache directly
struct file_bytes {
  const char* ptr{nullptr};
  size_t len{0};
  explicit file_bytes(const char* path)
  {
    B200_EXPECTS(path != nullptr, kInvalidInput, "path is NULL");
    int fd = open(path, O_RDONLY);
    B200_EXPECTS(fd >= 0, kInvalidInput, std::string("fopen (") + path + ") failure.");
    struct stat st {};
    bool ok = fstat(fd, &st) == 0 && S_ISREG(st.st_mode);
    if (ok && st.st_size > 0) {
      void* m = mmap(nullptr, static_cast<size_t>(st.st_size), PROT_READ, MAP_PRIVATE, fd, 0);
      ok      = m != MAP_FAILED;
      if (ok) {
        ptr = static_cast<const char*>(m);
        len = static_cast<size_t>(st.st_size);
        (void)madvise(m, len, MADV_SEQUENTIAL);
      }
    }
    close(fd);
    B200_EXPECTS(ok, kInvalidInput, std::string("could not read ") + path);
  }
  ~file_bytes()
  {
    if (ptr) munmap(const_cast<char*>(ptr), len);
  }
  file_bytes(file_bytes const&)            = delete;
  file_bytes& operator=(file_bytes const&) = delete;
  const char* begin() const { return ptr ? ptr : ""; }
  const char* end() const { return begin() + len; }
};

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }

// end of the line starting at p (position of '\n' or end)
inline const char* line_end(const char* p, const char* end)
{
  const void* nl = memchr(p, '\n', static_cast<size_t>(end - p));
  return nl ? static_cast<const char*>(nl) : end;
}

inline const char* skip_spaces(const char* p, const char* e)
{
  while (p < e && is_space(*p)) ++p;
  return p;
}

inline bool blank(const char* p, const char* e) { return skip_spaces(p, e) == e; }

template <typename T>
inline bool parse_number(const char*& p, const char* e, T& out)
{
  p = skip_spaces(p, e);
  if (p < e && *p == '+') ++p;  // from_chars takes no leading '+', fscanf does
  auto r = std::from_chars(p, e, out);
  if (r.ec != std::errc()) return false;
  p = r.ptr;
  return true;
}

struct parse_options {
  bool mtx{false};       // 1-based ids; mirrored entries when `mirror`
  bool mirror{false};
  bool skew{false};
  bool has_values{false};   // a value column is expected in the file
  bool want_values{false};  // ... and kept
  char delimiter{' '};
};


inline void* host_alloc(size_t bytes, bool& pinned)
{
  void* p = nullptr;
  if (posix_memalign(&p, 4096, std::max<size_t>(bytes, 4096)) != 0 || p == nullptr)
    throw error(kAllocError, "host allocation failed");
  // page-lock when a device is there; a host without one (the CPU test-suite) keeps pageable memory
  pinned = cudaHostRegister(p, std::max<size_t>(bytes, 4096), cudaHostRegisterDefault) == cudaSuccess;
  if (!pinned) (void)cudaGetLastError();
  return p;
}


inline std::string lower(std::string s)
{
  for (auto& c : s) c = static_cast<char>(tolower(static_cast<unsigned char>(c)));
  return s;
}

inline std::vector<std::string> tokens(const char* p, const char* e)
{
  std::vector<std::string> out;
  while (true) {
    p = skip_spaces(p, e);
    if (p >= e) break;
    const char* q = p;
    while (q < e && !is_space(*q)) ++q;
    out.emplace_back(p, q);
    p = q;
  }
  return out;
}

inline void check_types(int vertex_type, int weight_type, bool weights_optional)
{
  B200_EXPECTS(vertex_type == B200_INT32 || vertex_type == B200_INT64, kInvalidInput,
               "vertex_type must be B200_INT32 or B200_INT64");
  B200_EXPECTS(weight_type == B200_FLOAT32 || weight_type == B200_FLOAT64 ||
                 (weights_optional && weight_type < 0),
               kInvalidInput, "weight_type must be B200_FLOAT32 or B200_FLOAT64");
}

template <typename F>
int guarded(F&& f)
{
  try {
    f();
    return kSuccess;
  } catch (error const& e) {
    set_last_error(e.what());
    return e.code;
  } catch (std::exception const& e) {
    set_last_error(e.what());
    return kUnknownError;
  }
}


// what the banner and the size line of a Matrix-Market file say, and where the entries start
struct mtx_header {
  int64_t rows{-1}, cols{-1}, nnz{-1};
  bool symmetric{false};   // banner says "symmetric"
  bool mirror{false};      // anything but "general": off-diagonal entries are mirrored
  bool pattern{false};
  const char* body{nullptr};
  int64_t body_line{0};    // 0-based line number of `body`
};

inline mtx_header read_mtx_header(file_bytes const& file)
{
  mtx_header h;
  const char* p   = file.begin();
  const char* end = file.end();
  // banner (mmio mm_read_banner): %%MatrixMarket matrix coordinate <field> <symmetry>
  const char* e = line_end(p, end);
  auto banner   = tokens(p, e);
  B200_EXPECTS(banner.size() >= 5 && lower(banner[0]) == "%%matrixmarket", kInvalidInput,
               "could not read Matrix Market file properties.");
  std::string object = lower(banner[1]), format = lower(banner[2]), field = lower(banner[3]),
              symmetry = lower(banner[4]);
  bool known_field = field == "pattern" || field == "real" || field == "integer" ||
                     field == "complex";
  bool known_sym   = symmetry == "general" || symmetry == "symmetric" ||
                   symmetry == "skew-symmetric" || symmetry == "hermitian";
  B200_EXPECTS(known_field && known_sym, kInvalidInput,
               "could not read Matrix Market file properties.");
  // read_edgelist_from_matrix_market_file :289
  B200_EXPECTS(object == "matrix" && format == "coordinate" && field != "complex" &&
                 symmetry != "skew-symmetric",
               kInvalidInput, "invalid Matrix Market file properties.");
  int64_t line_no = 1;
  p               = e < end ? e + 1 : end;
  // comments and blank lines, then the size line (mm_read_mtx_crd_size)
  while (p < end) {
    e = line_end(p, end);
    if (!blank(p, e) && *skip_spaces(p, e) != '%') {
      const char* q = p;
      bool ok = parse_number(q, e, h.rows) && parse_number(q, e, h.cols) && parse_number(q, e, h.nnz);
      B200_EXPECTS(ok && h.rows >= 0 && h.cols >= 0 && h.nnz >= 0, kInvalidInput,
                   "could not read Matrix Market file properties.");
      p = e < end ? e + 1 : end;
      ++line_no;
      break;
    }
    p = e < end ? e + 1 : end;
    ++line_no;
  }
  B200_EXPECTS(h.nnz >= 0, kInvalidInput, "could not read Matrix Market file properties.");
  h.symmetric = symmetry == "symmetric";
  h.mirror    = symmetry != "general";
  h.pattern   = field == "pattern";
  h.body      = p;
  h.body_line = line_no;
  return h;
}

inline void check_delimiter(char delimiter)
{
  B200_EXPECTS(delimiter != '\n' && delimiter != '\0' && delimiter != '.' && delimiter != '-' &&
                 delimiter != '+' && !(delimiter >= '0' && delimiter <= '9'),
               kInvalidInput, "invalid delimiter");
}

// start of the entries of a delimited file after `skip_rows` lines
inline const char* csv_body(file_bytes const& file, int skip_rows, int64_t& line_no)
{
  const char* p   = file.begin();
  const char* end = file.end();
  for (line_no = 0; line_no < skip_rows && p < end; ++line_no) {
    const char* e = line_end(p, end);
    p             = e < end ? e + 1 : end;
  }
  return p;
}

// the host parser (ingest.cu): text [begin, end) -> COO in host memory, `threads_arg` host threads
b200_host_edgelist* parse_body(const char* begin, const char* end, int64_t first_line,
                               parse_options const& o, int vertex_type, int weight_type,
                               int threads_arg);

// the size-line checks of a parsed Matrix-Market body (stricter than the reference's fscanf loop)
inline void check_mtx_counts(mtx_header const& h, int64_t file_entries, int64_t max_id)
{
  if (file_entries != h.nnz)
    throw error(kInvalidInput, "could not read matrix data (" + std::to_string(file_entries) +
                                 " entries, size line says " + std::to_string(h.nnz) + ")");
  if (max_id + 1 > std::max(h.rows, h.cols))
    throw error(kInvalidInput, "could not read matrix data (index beyond the size line)");
}

}  // namespace ingest
}  // namespace b200
