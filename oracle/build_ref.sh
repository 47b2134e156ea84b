#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Builds
#   oracle/libcentrality_oracle.so      (our CPU restatement, oracle/centrality_oracle.cpp)
#   oracle/_ref/libcugraph_ref_oracle.so (the reference's own host oracles - centrality, HITS,
#                                         graph structure, Matrix-Market reader - compiled from
#                                         the sources where they lie under /root/reference)
# The second target is skipped when /root/reference is absent (GPU box): the prebuilt .so
# travels with the snapshot. No reference source text is written into the repository: the
# extracted function bodies live in a mktemp directory that is removed on exit.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ref="${CUGRAPH_REFERENCE_ROOT:-/root/reference}"
cxx="${CXX:-g++}"
# no -march=native: the prebuilt _ref .so must run on whatever host CPU the GPU box has
flags="-O3 -std=c++17 -fPIC -shared -fno-fast-math -ffp-contract=off"

$cxx $flags -pthread -o "$here/libcentrality_oracle.so" "$here/centrality_oracle.cpp" "$here/inputs_oracle.cpp"

if [ -d "$ref/cpp/tests" ]; then
  tmp="$(mktemp -d)"
  trap 'rm -rf "$tmp"' EXIT
  # function-template line ranges (template<...> line through the closing brace)
  sed -n '42,130p' "$ref/cpp/tests/link_analysis/pagerank_test.cpp"            > "$tmp/ref_pagerank.inc"
  sed -n '41,93p'  "$ref/cpp/tests/centrality/katz_centrality_test.cpp"        > "$tmp/ref_katz.inc"
  sed -n '41,96p'  "$ref/cpp/tests/centrality/eigenvector_centrality_test.cpp" > "$tmp/ref_eigenvector.inc"
  sed -n '46,127p' "$ref/cpp/tests/link_analysis/hits_test.cpp"                > "$tmp/ref_hits.inc"
  sed -n '53,218p' "$ref/cpp/tests/utilities/matrix_market_file_utilities.cu"  > "$tmp/ref_mm.inc"
  sed -n '38,69p'  "$ref/cpp/tests/structure/graph_test.cpp"                   > "$tmp/ref_graph.inc"
  sed -n '36,57p'  "$ref/cpp/tests/structure/degree_test.cpp"                  > "$tmp/ref_degree.inc"
  sed -n '37,59p'  "$ref/cpp/tests/structure/weight_sum_test.cpp"              > "$tmp/ref_weight_sum.inc"
  grep -q 'graph_reference('                      "$tmp/ref_graph.inc"
  grep -q 'void degree_reference('                "$tmp/ref_degree.inc"
  grep -q 'void weight_sum_reference('            "$tmp/ref_weight_sum.inc"
  grep -q 'int mm_properties('                    "$tmp/ref_mm.inc"
  grep -q 'int mm_to_coo('                        "$tmp/ref_mm.inc"
  grep -q 'hits_reference('                       "$tmp/ref_hits.inc"
  grep -q 'void pagerank_reference'               "$tmp/ref_pagerank.inc"
  grep -q 'void katz_centrality_reference'        "$tmp/ref_katz.inc"
  grep -q 'void eigenvector_centrality_reference' "$tmp/ref_eigenvector.inc"
  mkdir -p "$here/_ref"
  ${CC:-gcc} -O2 -fPIC -w -c "$ref/thirdparty/mmio/mmio.c" -o "$tmp/mmio.o"
  $cxx $flags -I"$tmp" -I"$ref/thirdparty/mmio" -o "$here/_ref/libcugraph_ref_oracle.so" \
    "$here/ref_shim.cpp" "$here/ref_mtx_shim.cpp" "$here/ref_structure_shim.cpp" "$tmp/mmio.o"
  echo "built oracle/_ref/libcugraph_ref_oracle.so from $ref"
else
  echo "reference tree not present; keeping prebuilt oracle/_ref (if any)"
fi
