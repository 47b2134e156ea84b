/*
 * cugraph_c/resource_handle.h - drop-in for cpp/include/cugraph_c/resource_handle.h.
 * The reference wraps a raft::handle_t*; raft is not part of this library, so the argument of
 * cugraph_create_resource_handle is either NULL (single GPU, private stream) or a
 * b200_handle_t* made by b200_handle_create / b200_handle_create_mg (include/b200_cugraph.h),
 * which carries the stream and the NCCL world / row / column communicators a raft handle
 * with "comm_p_row" / "comm_p_col" sub-communicators would carry.
 */
#pragma once

#include <cugraph_c/error.h>

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum bool_ { FALSE = 0, TRUE = 1 } bool_t;

typedef int8_t byte_t;

typedef enum data_type_id_ { INT32 = 0, INT64, FLOAT32, FLOAT64, NTYPES } data_type_id_t;

typedef struct cugraph_resource_handle_ {
  int32_t align_;
} cugraph_resource_handle_t;

cugraph_resource_handle_t* cugraph_create_resource_handle(void* raft_handle);
int cugraph_resource_handle_get_rank(const cugraph_resource_handle_t* handle);
void cugraph_free_resource_handle(cugraph_resource_handle_t* handle);

#ifdef __cplusplus
}
#endif
