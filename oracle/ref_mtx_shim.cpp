// TEST INFRASTRUCTURE ONLY. C entry points around the reference's own Matrix-Market reader:
// mm_properties and mm_to_coo of /root/reference/cpp/tests/utilities/
// matrix_market_file_utilities.cu:53-218, compiled IN PLACE (build_ref.sh extracts the two
// function templates into a temporary ref_mm.inc; nothing of them is stored in this repository)
// together with /root/reference/thirdparty/mmio/mmio.c.  The call sequence below is the one of
// read_edgelist_from_matrix_market_file (:283-304).  Used by tests/ to pin the oracle's
// restatement (oracle/graph_oracle.py read_mtx) and the library's reader (csrc/ingest.cu).
#include <cstdint>
#include <cstdio>
#include <vector>

extern "C" {
#include "mmio.h"
}

namespace cugraph {
namespace test {
#include "ref_mm.inc"
}  // namespace test
}  // namespace cugraph

extern "C" {

// returns 0 and fills the outputs, or a negative code: -1 fopen, -2 properties, -3 rejected type
long ref_mtx_properties(const char* path, int64_t* rows, int64_t* cols, int64_t* nnz,
                        int* is_symmetric, int* is_pattern)
{
  FILE* f = fopen(path, "r");
  if (f == nullptr) return -1;
  MM_typecode mc{};
  size_t m{}, k{}, n{};
  int ret = cugraph::test::mm_properties<size_t>(f, 1, &mc, &m, &k, &n);
  fclose(f);
  if (ret != 0) return -2;
  if (!(mm_is_matrix(mc) && mm_is_coordinate(mc) && !mm_is_complex(mc) && !mm_is_skew(mc)))
    return -3;
  *rows         = static_cast<int64_t>(m);
  *cols         = static_cast<int64_t>(k);
  *nnz          = static_cast<int64_t>(n);
  *is_symmetric = mm_is_symmetric(mc) ? 1 : 0;
  *is_pattern   = mm_is_pattern(mc) ? 1 : 0;
  return 0;
}

#define REF_MTX_TO_COO(SFX, W)                                                                   \
  long ref_mtx_to_coo_##SFX(const char* path, int64_t nnz, int32_t* rows, int32_t* cols, W* vals) \
  {                                                                                              \
    FILE* f = fopen(path, "r");                                                                  \
    if (f == nullptr) return -1;                                                                 \
    int ret = cugraph::test::mm_to_coo<int32_t, W>(f, 1, static_cast<int32_t>(nnz), rows, cols,  \
                                                   vals, static_cast<W*>(nullptr));              \
    fclose(f);                                                                                   \
    return ret;                                                                                  \
  }
REF_MTX_TO_COO(f32, float)
REF_MTX_TO_COO(f64, double)

}  // extern "C"
