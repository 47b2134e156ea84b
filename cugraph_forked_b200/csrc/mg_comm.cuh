// NCCL plumbing of the multi-GPU path: host-scalar gathers, variable all-to-all, allgatherv.
// What raft::comms + cugraph's shuffle_comm.cuh / host_scalar_comm.hpp are for the reference
// (cpp/include/cugraph/utilities/shuffle_comm.cuh, host_scalar_comm.hpp).  Everything runs on
// the handle's stream; element counts travel as bytes so one code path serves every type.
#pragma once

#include <numeric>
#include <vector>

#include "common.cuh"

namespace b200 {

// every rank contributes k int64 values; returns the comm_size * k values in rank order
inline std::vector<int64_t> allgather_host(b200_handle* h, ncclComm_t comm, int comm_size,
                                           std::vector<int64_t> const& mine)
{
  cudaStream_t s = h->stream;
  size_t k       = mine.size();
  std::vector<int64_t> all(k * comm_size);
  if (k == 0) return all;
  if (comm_size == 1) return mine;
  dbuf<int64_t> d_in(k, s), d_out(k * comm_size, s);
  CUDA_TRY(cudaMemcpyAsync(d_in.data(), mine.data(), k * 8, cudaMemcpyHostToDevice, s));
  NCCL_TRY(ncclAllGather(d_in.data(), d_out.data(), k, ncclInt64, comm, s));
  CUDA_TRY(cudaMemcpyAsync(all.data(), d_out.data(), k * comm_size * 8, cudaMemcpyDeviceToHost, s));
  CUDA_TRY(cudaStreamSynchronize(s));
  return all;
}

// send_counts[p] elements go to rank p (send buffer grouped by destination, in rank order);
// returns recv_counts[p] = elements rank p sends to this rank
inline std::vector<size_t> exchange_counts(b200_handle* h, ncclComm_t comm, int comm_size,
                                           int comm_rank, std::vector<size_t> const& send_counts)
{
  std::vector<int64_t> mine(send_counts.begin(), send_counts.end());
  std::vector<int64_t> all = allgather_host(h, comm, comm_size, mine);
  std::vector<size_t> recv(comm_size);
  for (int p = 0; p < comm_size; ++p) recv[p] = static_cast<size_t>(all[p * comm_size + comm_rank]);
  return recv;
}

// variable all-to-all of `elem_bytes`-sized elements (grouped ncclSend / ncclRecv; the piece a
// rank keeps for itself is a device copy)
inline void alltoallv(b200_handle* h, ncclComm_t comm, int comm_size, int comm_rank,
                      const void* send, std::vector<size_t> const& send_counts, void* recv,
                      std::vector<size_t> const& recv_counts, size_t elem_bytes)
{
  cudaStream_t s   = h->stream;
  const char* sbuf = static_cast<const char*>(send);
  char* rbuf       = static_cast<char*>(recv);
  std::vector<size_t> sdis(comm_size + 1, 0), rdis(comm_size + 1, 0);
  for (int p = 0; p < comm_size; ++p) {
    sdis[p + 1] = sdis[p] + send_counts[p];
    rdis[p + 1] = rdis[p] + recv_counts[p];
  }
  if (send_counts[comm_rank])
    CUDA_TRY(cudaMemcpyAsync(rbuf + rdis[comm_rank] * elem_bytes, sbuf + sdis[comm_rank] * elem_bytes,
                             send_counts[comm_rank] * elem_bytes, cudaMemcpyDeviceToDevice, s));
  if (comm_size == 1) return;
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < comm_size; ++p) {
    if (p == comm_rank) continue;
    if (send_counts[p])
      NCCL_TRY(ncclSend(sbuf + sdis[p] * elem_bytes, send_counts[p] * elem_bytes, ncclInt8, p, comm, s));
    if (recv_counts[p])
      NCCL_TRY(ncclRecv(rbuf + rdis[p] * elem_bytes, recv_counts[p] * elem_bytes, ncclInt8, p, comm, s));
  }
  NCCL_TRY(ncclGroupEnd());
}

// allgatherv: rank p contributes counts[p] elements, everyone ends with the concatenation in
// rank order.  Point-to-point form (every rank sends its piece to every peer inside one group):
// on NVSwitch each pair has its own full-bandwidth path, and for the 2- and 4-rank row
// communicators of an 8-GPU box this measured faster than one broadcast per root.
inline void allgatherv(cudaStream_t s, ncclComm_t comm, int comm_size, int comm_rank,
                       const void* send, void* recv, std::vector<size_t> const& counts,
                       size_t elem_bytes)
{
  char* rbuf = static_cast<char*>(recv);
  std::vector<size_t> dis(comm_size + 1, 0);
  for (int p = 0; p < comm_size; ++p) dis[p + 1] = dis[p] + counts[p];
  if (counts[comm_rank] && send != rbuf + dis[comm_rank] * elem_bytes)
    CUDA_TRY(cudaMemcpyAsync(rbuf + dis[comm_rank] * elem_bytes, send, counts[comm_rank] * elem_bytes,
                             cudaMemcpyDeviceToDevice, s));
  if (comm_size == 1) return;
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < comm_size; ++p) {
    if (p == comm_rank) continue;
    if (counts[comm_rank])
      NCCL_TRY(ncclSend(send, counts[comm_rank] * elem_bytes, ncclInt8, p, comm, s));
    if (counts[p])
      NCCL_TRY(ncclRecv(rbuf + dis[p] * elem_bytes, counts[p] * elem_bytes, ncclInt8, p, comm, s));
  }
  NCCL_TRY(ncclGroupEnd());
}

}  // namespace b200
