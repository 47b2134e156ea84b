// C++ client of the rows after the hot path: the file reader (b200_read_matrix_market, the
// stand-in for read_edgelist_from_matrix_market_file) feeding the cugraph::hits shim of
// include/cugraph_b200/algorithms.hpp, on the golden graph of cpp/tests/c_api/hits_test.c
// (test_hits: 6 vertices, 8 edges, epsilon 1e-4, 20 iterations, normalize = false).
// Build: g++ -std=c++17 -I include -I /usr/local/cuda/include hits_file_golden.cpp
//            -L cugraph_forked_b200 -lcugraph_b200 -L /usr/local/cuda/lib64 -lcudart
#include <cugraph_b200/algorithms.hpp>

#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <tuple>
#include <vector>

static bool nearly_equal(double a, double b, double eps = 1e-3)
{
  return std::fabs(a - b) <= std::max(std::fabs(a), std::fabs(b)) * eps;
}

int main()
{
  // the golden graph as a Matrix-Market file (1-based ids)
  std::string path = std::string(getenv("TMPDIR") ? getenv("TMPDIR") : "/tmp") + "/b200_hits_golden.mtx";
  FILE* f          = fopen(path.c_str(), "w");
  if (f == nullptr) return 1;
  fprintf(f,
          "%%%%MatrixMarket matrix coordinate real general\n%% hits_test.c test_hits\n6 6 8\n"
          "1 2 0.1\n2 4 2.1\n2 5 1.1\n3 1 5.1\n3 2 3.1\n3 4 4.1\n4 6 7.2\n5 6 3.2\n");
  fclose(f);

  b200_host_edgelist_t* el = nullptr;
  cugraph::b200::check(b200_read_matrix_market(path.c_str(), B200_INT32, B200_FLOAT32, 0, &el), "read");
  remove(path.c_str());
  b200_host_edgelist_info_t info{};
  b200_host_edgelist_info(el, &info);
  if (info.num_rows != 6 || info.num_edges != 8 || info.is_symmetric || info.is_pattern) {
    std::printf("reader mismatch\n");
    return 1;
  }
  auto const* h_src = static_cast<const int32_t*>(b200_host_edgelist_src(el));
  auto const* h_wgt = static_cast<const float*>(b200_host_edgelist_weights(el));
  if (h_src[3] != 2 || h_wgt[6] != 7.2f) { std::printf("reader values\n"); return 1; }

  size_t n = static_cast<size_t>(info.num_edges);
  int32_t *d_src = nullptr, *d_dst = nullptr;
  float* d_wgt = nullptr;
  cudaMalloc(reinterpret_cast<void**>(&d_src), n * 4);
  cudaMalloc(reinterpret_cast<void**>(&d_dst), n * 4);
  cudaMalloc(reinterpret_cast<void**>(&d_wgt), n * 4);
  cudaMemcpy(d_src, b200_host_edgelist_src(el), n * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_dst, b200_host_edgelist_dst(el), n * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_wgt, b200_host_edgelist_weights(el), n * 4, cudaMemcpyHostToDevice);
  b200_host_edgelist_free(el);

  cugraph::b200::handle_t handle;
  b200_graph_t* g = nullptr;
  cugraph::b200::check(b200_graph_create_sg(handle.get(), d_src, d_dst, d_wgt, n, B200_INT32,
                                            B200_FLOAT32, nullptr, 0, 1, 1, &g),
                       "graph");
  cugraph::graph_view_t<int32_t, int32_t, float, true, false> view(g);
  b200_graph_view_t raw;
  b200_graph_view(g, &raw);
  std::vector<int32_t> labels(6);
  cudaMemcpy(labels.data(), raw.renumber_map, 6 * 4, cudaMemcpyDeviceToHost);

  float *d_hubs = nullptr, *d_auth = nullptr;
  cudaMalloc(reinterpret_cast<void**>(&d_hubs), 6 * sizeof(float));
  cudaMalloc(reinterpret_cast<void**>(&d_auth), 6 * sizeof(float));
  auto [diff, iterations] = cugraph::hits<int32_t, int32_t, float, false>(
    handle, view, d_hubs, d_auth, 1e-4f, 20, false, false, false);
  std::vector<float> hubs(6), auth(6);
  cudaMemcpy(hubs.data(), d_hubs, 6 * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(auth.data(), d_auth, 6 * 4, cudaMemcpyDeviceToHost);
  double hubs_expect[6] = {0.347296, 0.532089, 1, 0.00003608, 0.00003608, 0};
  double auth_expect[6] = {0.652703, 0.879385, 0, 1, 0.347296, 0.00009136};
  for (int i = 0; i < 6; ++i) {
    if (!nearly_equal(hubs[i], hubs_expect[labels[i]]) || !nearly_equal(auth[i], auth_expect[labels[i]])) {
      std::printf("hits mismatch at %d\n", i);
      return 1;
    }
  }
  if (iterations == 0 || iterations > 20 || !(diff >= 0)) { std::printf("hits stats\n"); return 1; }

  b200_graph_free(g);
  cudaFree(d_src); cudaFree(d_dst); cudaFree(d_wgt); cudaFree(d_hubs); cudaFree(d_auth);
  std::printf("HITS_FILE_GOLDEN_OK\n");
  return 0;
}
