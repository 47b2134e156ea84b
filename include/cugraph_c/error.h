/*
 * cugraph_c/error.h - drop-in for the reference's cpp/include/cugraph_c/error.h (same enum
 * values, same opaque error object, same two calls); implemented by libcugraph_b200.so.
 */
#pragma once

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cugraph_error_code_ {
  CUGRAPH_SUCCESS = 0,
  CUGRAPH_UNKNOWN_ERROR,
  CUGRAPH_INVALID_HANDLE,
  CUGRAPH_ALLOC_ERROR,
  CUGRAPH_INVALID_INPUT,
  CUGRAPH_NOT_IMPLEMENTED,
  CUGRAPH_UNSUPPORTED_TYPE_COMBINATION
} cugraph_error_code_t;

typedef struct cugraph_error_ {
  int32_t align_;
} cugraph_error_t;

/* message of an error object returned through a `cugraph_error_t**` argument */
const char* cugraph_error_message(const cugraph_error_t* error);
void cugraph_error_free(cugraph_error_t* error);

#ifdef __cplusplus
}
#endif
