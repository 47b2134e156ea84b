/*
 * cugraph_c/array.h - drop-in for cpp/include/cugraph_c/array.h:52-326: type-erased device /
 * host arrays and non-owning views.  Same names, argument order and ownership rules.
 */
#pragma once

#include <cugraph_c/resource_handle.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t align_;
} cugraph_type_erased_device_array_t;

typedef struct {
  int32_t align_;
} cugraph_type_erased_device_array_view_t;

typedef struct {
  int32_t align_;
} cugraph_type_erased_host_array_t;

typedef struct {
  int32_t align_;
} cugraph_type_erased_host_array_view_t;

cugraph_error_code_t cugraph_type_erased_device_array_create(
  const cugraph_resource_handle_t* handle,
  size_t n_elems,
  data_type_id_t dtype,
  cugraph_type_erased_device_array_t** array,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_type_erased_device_array_create_from_view(
  const cugraph_resource_handle_t* handle,
  const cugraph_type_erased_device_array_view_t* view,
  cugraph_type_erased_device_array_t** array,
  cugraph_error_t** error);

void cugraph_type_erased_device_array_free(cugraph_type_erased_device_array_t* p);

cugraph_type_erased_device_array_view_t* cugraph_type_erased_device_array_view(
  cugraph_type_erased_device_array_t* array);

cugraph_error_code_t cugraph_type_erased_device_array_view_as_type(
  cugraph_type_erased_device_array_t* array,
  data_type_id_t dtype,
  cugraph_type_erased_device_array_view_t** result_view,
  cugraph_error_t** error);

cugraph_type_erased_device_array_view_t* cugraph_type_erased_device_array_view_create(
  void* pointer, size_t n_elems, data_type_id_t dtype);

void cugraph_type_erased_device_array_view_free(cugraph_type_erased_device_array_view_t* p);

size_t cugraph_type_erased_device_array_view_size(const cugraph_type_erased_device_array_view_t* p);

data_type_id_t cugraph_type_erased_device_array_view_type(
  const cugraph_type_erased_device_array_view_t* p);

const void* cugraph_type_erased_device_array_view_pointer(
  const cugraph_type_erased_device_array_view_t* p);

cugraph_error_code_t cugraph_type_erased_host_array_create(const cugraph_resource_handle_t* handle,
                                                           size_t n_elems,
                                                           data_type_id_t dtype,
                                                           cugraph_type_erased_host_array_t** array,
                                                           cugraph_error_t** error);

void cugraph_type_erased_host_array_free(cugraph_type_erased_host_array_t* p);

cugraph_type_erased_host_array_view_t* cugraph_type_erased_host_array_view(
  cugraph_type_erased_host_array_t* array);

cugraph_type_erased_host_array_view_t* cugraph_type_erased_host_array_view_create(
  void* pointer, size_t n_elems, data_type_id_t dtype);

void cugraph_type_erased_host_array_view_free(cugraph_type_erased_host_array_view_t* p);

size_t cugraph_type_erased_host_array_size(const cugraph_type_erased_host_array_view_t* p);

data_type_id_t cugraph_type_erased_host_array_type(const cugraph_type_erased_host_array_view_t* p);

void* cugraph_type_erased_host_array_pointer(const cugraph_type_erased_host_array_view_t* p);

cugraph_error_code_t cugraph_type_erased_host_array_view_copy(
  const cugraph_resource_handle_t* handle,
  cugraph_type_erased_host_array_view_t* dst,
  const cugraph_type_erased_host_array_view_t* src,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_type_erased_device_array_view_copy_from_host(
  const cugraph_resource_handle_t* handle,
  cugraph_type_erased_device_array_view_t* dst,
  const byte_t* h_src,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_type_erased_device_array_view_copy_to_host(
  const cugraph_resource_handle_t* handle,
  byte_t* h_dst,
  const cugraph_type_erased_device_array_view_t* src,
  cugraph_error_t** error);

cugraph_error_code_t cugraph_type_erased_device_array_view_copy(
  const cugraph_resource_handle_t* handle,
  cugraph_type_erased_device_array_view_t* dst,
  const cugraph_type_erased_device_array_view_t* src,
  cugraph_error_t** error);

#ifdef __cplusplus
}
#endif
