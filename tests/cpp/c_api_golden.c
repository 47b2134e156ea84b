/* Pure-C client of the cugraph_c ABI, shaped like the reference's own C tests
 * (cpp/tests/c_api/pagerank_test.c:35-209 generic_pagerank_test, :211-228 test_pagerank):
 * host COO -> type-erased device arrays -> cugraph_sg_graph_create -> cugraph_pagerank ->
 * copy (vertices, values) back and compare with the golden vector at relative 1e-3
 * (cpp/tests/c_api/test_utils.cpp:21-27).  Exit code 0 = pass.
 * Build: gcc -std=c99 -I include c_api_golden.c -L cugraph_forked_b200 -lcugraph_b200 */
#include <cugraph_c/algorithms.h>

#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#define CHECK(cond, msg)                                     \
  do {                                                       \
    if (!(cond)) {                                           \
      fprintf(stderr, "FAILED: %s (%s)\n", msg,              \
              err ? cugraph_error_message(err) : "no error"); \
      return 1;                                              \
    }                                                        \
  } while (0)

static int nearly_equal(double a, double b, double eps)
{
  double m = fabs(a) > fabs(b) ? fabs(a) : fabs(b);
  return fabs(a - b) <= m * eps;
}

static int to_device(const cugraph_resource_handle_t* h, const void* host, size_t n,
                     data_type_id_t t, cugraph_type_erased_device_array_t** arr,
                     cugraph_type_erased_device_array_view_t** view, cugraph_error_t** err)
{
  if (cugraph_type_erased_device_array_create(h, n, t, arr, err) != CUGRAPH_SUCCESS) return 1;
  *view = cugraph_type_erased_device_array_view(*arr);
  return cugraph_type_erased_device_array_view_copy_from_host(h, *view, (const byte_t*)host, err) !=
         CUGRAPH_SUCCESS;
}

static int run(bool_t store_transposed)
{
  int32_t h_src[] = {0, 1, 1, 2, 2, 2, 3, 4};
  int32_t h_dst[] = {1, 3, 4, 0, 1, 3, 5, 5};
  float h_wgt[]   = {0.1f, 2.1f, 1.1f, 5.1f, 3.1f, 4.1f, 7.2f, 3.2f};
  float expect[]  = {0.0915528f, 0.168382f, 0.0656831f, 0.191468f, 0.120677f, 0.362237f};
  size_t ne = 8, nv = 6;
  cugraph_error_t* err = NULL;
  cugraph_resource_handle_t* h = cugraph_create_resource_handle(NULL);
  CHECK(h != NULL, "resource handle");
  cugraph_type_erased_device_array_t *src, *dst, *wgt;
  cugraph_type_erased_device_array_view_t *vsrc, *vdst, *vwgt;
  CHECK(!to_device(h, h_src, ne, INT32, &src, &vsrc, &err), "src");
  CHECK(!to_device(h, h_dst, ne, INT32, &dst, &vdst, &err), "dst");
  CHECK(!to_device(h, h_wgt, ne, FLOAT32, &wgt, &vwgt, &err), "wgt");
  cugraph_graph_properties_t props;
  props.is_symmetric  = FALSE;
  props.is_multigraph = FALSE;
  cugraph_graph_t* g = NULL;
  CHECK(cugraph_sg_graph_create(h, &props, vsrc, vdst, vwgt, NULL, NULL, store_transposed, FALSE,
                                FALSE, &g, &err) == CUGRAPH_SUCCESS,
        "graph create");
  cugraph_centrality_result_t* res = NULL;
  CHECK(cugraph_pagerank(h, g, NULL, NULL, NULL, NULL, 0.95, 0.0001, 20, FALSE, &res, &err) ==
          CUGRAPH_SUCCESS,
        "pagerank");
  cugraph_type_erased_device_array_view_t* verts = cugraph_centrality_result_get_vertices(res);
  cugraph_type_erased_device_array_view_t* vals  = cugraph_centrality_result_get_values(res);
  CHECK(cugraph_type_erased_device_array_view_size(verts) == nv, "result size");
  int32_t h_v[6];
  float h_r[6];
  CHECK(cugraph_type_erased_device_array_view_copy_to_host(h, (byte_t*)h_v, verts, &err) ==
          CUGRAPH_SUCCESS,
        "copy vertices");
  CHECK(cugraph_type_erased_device_array_view_copy_to_host(h, (byte_t*)h_r, vals, &err) ==
          CUGRAPH_SUCCESS,
        "copy values");
  for (size_t i = 0; i < nv; ++i)
    CHECK(nearly_equal(expect[h_v[i]], h_r[i], 0.001), "pagerank value");
  /* error convention: size mismatch -> CUGRAPH_INVALID_INPUT with a message, *graph == NULL */
  cugraph_type_erased_device_array_view_t* half =
    cugraph_type_erased_device_array_view_create((void*)cugraph_type_erased_device_array_view_pointer(vdst),
                                                 4, INT32);
  cugraph_graph_t* g2 = NULL;
  cugraph_error_t* err2 = NULL;
  cugraph_error_code_t rc = cugraph_sg_graph_create(h, &props, vsrc, half, NULL, NULL, NULL, TRUE,
                                                    FALSE, FALSE, &g2, &err2);
  if (rc != CUGRAPH_INVALID_INPUT || g2 != NULL || err2 == NULL) {
    fprintf(stderr, "FAILED: error convention\n");
    return 1;
  }
  cugraph_error_free(err2);
  cugraph_type_erased_device_array_view_free(half);
  cugraph_type_erased_device_array_view_free(verts);
  cugraph_type_erased_device_array_view_free(vals);
  cugraph_centrality_result_free(res);
  cugraph_sg_graph_free(g);
  cugraph_type_erased_device_array_view_free(vsrc);
  cugraph_type_erased_device_array_view_free(vdst);
  cugraph_type_erased_device_array_view_free(vwgt);
  cugraph_type_erased_device_array_free(src);
  cugraph_type_erased_device_array_free(dst);
  cugraph_type_erased_device_array_free(wgt);
  cugraph_free_resource_handle(h);
  return 0;
}

int main(void)
{
  if (run(TRUE)) return 1;
  if (run(FALSE)) return 1; /* cugraph_pagerank transposes the graph in place */
  printf("C_API_GOLDEN_OK\n");
  return 0;
}
