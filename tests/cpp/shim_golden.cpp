// C++ client of include/cugraph_b200/algorithms.hpp: the reference's cugraph::pagerank /
// katz_centrality signatures over the raw ABI, on the golden graph of
// cpp/tests/c_api/pagerank_test.c:211-228 and katz_test.c:103-129.
// Build: g++ -std=c++17 -I include -I /usr/local/cuda/include shim_golden.cpp
//            -L cugraph_forked_b200 -lcugraph_b200 -L /usr/local/cuda/lib64 -lcudart
#include <cugraph_b200/algorithms.hpp>

#include <cuda_runtime_api.h>

#include <cmath>
#include <cstdio>
#include <vector>

template <typename T>
static T* to_device(std::vector<T> const& v)
{
  T* d = nullptr;
  cudaMalloc(reinterpret_cast<void**>(&d), v.size() * sizeof(T));
  cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice);
  return d;
}

static bool nearly_equal(double a, double b, double eps = 1e-3)
{
  return std::fabs(a - b) <= std::max(std::fabs(a), std::fabs(b)) * eps;
}

int main()
{
  std::vector<int32_t> src{0, 1, 1, 2, 2, 2, 3, 4}, dst{1, 3, 4, 0, 1, 3, 5, 5};
  std::vector<float> wgt{0.1f, 2.1f, 1.1f, 5.1f, 3.1f, 4.1f, 7.2f, 3.2f};
  std::vector<float> pr_expect{0.0915528f, 0.168382f, 0.0656831f, 0.191468f, 0.120677f, 0.362237f};
  std::vector<float> katz_expect{0.410614f, 0.403211f, 0.390689f, 0.415175f, 0.395125f, 0.433226f};
  cugraph::b200::handle_t handle;
  int32_t *d_src = to_device(src), *d_dst = to_device(dst);
  float* d_wgt = to_device(wgt);
  b200_graph_t* g = nullptr;
  cugraph::b200::check(b200_graph_create_sg(handle.get(), d_src, d_dst, d_wgt, src.size(), B200_INT32,
                                            B200_FLOAT32, nullptr, 0, 1, 1, &g),
                       "graph");
  cugraph::graph_view_t<int32_t, int32_t, float, true, false> view(g);
  if (view.number_of_vertices() != 6 || view.number_of_edges() != 8) return 1;
  b200_graph_view_t raw;
  b200_graph_view(g, &raw);
  std::vector<int32_t> labels(6);
  cudaMemcpy(labels.data(), raw.renumber_map, 6 * 4, cudaMemcpyDeviceToHost);
  float* d_out = nullptr;
  cudaMalloc(reinterpret_cast<void**>(&d_out), 6 * sizeof(float));
  std::vector<float> out(6);

  cugraph::pagerank<int32_t, int32_t, float, float, false>(handle, view, std::nullopt, std::nullopt,
                                                           std::nullopt, std::nullopt, d_out, 0.95f,
                                                           1e-4f, 20);
  cudaMemcpy(out.data(), d_out, 6 * 4, cudaMemcpyDeviceToHost);
  for (int i = 0; i < 6; ++i)
    if (!nearly_equal(out[i], pr_expect[labels[i]])) { std::printf("pagerank mismatch\n"); return 1; }

  cugraph::katz_centrality<int32_t, int32_t, float, float, false>(
    handle, view, nullptr, d_out, 0.01f, 1.0f, 1e-6f, 1000, false, true);
  cudaMemcpy(out.data(), d_out, 6 * 4, cudaMemcpyDeviceToHost);
  for (int i = 0; i < 6; ++i)
    if (!nearly_equal(out[i], katz_expect[labels[i]])) { std::printf("katz mismatch\n"); return 1; }

  // error behaviour: the reference throws on a non-converging run (pagerank_impl.cuh:289)
  bool threw = false;
  try {
    cugraph::pagerank<int32_t, int32_t, float, float, false>(handle, view, std::nullopt, std::nullopt,
                                                             std::nullopt, std::nullopt, d_out, 0.95f,
                                                             1e-12f, 2);
  } catch (cugraph::logic_error const& e) {
    threw = std::string(e.what()).find("failed to converge") != std::string::npos;
  }
  if (!threw) { std::printf("no convergence error\n"); return 1; }
  b200_graph_free(g);
  cudaFree(d_src); cudaFree(d_dst); cudaFree(d_wgt); cudaFree(d_out);
  std::printf("SHIM_GOLDEN_OK\n");
  return 0;
}
