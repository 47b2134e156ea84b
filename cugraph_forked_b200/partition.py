"""Host-side mirror of the reference's 2D partition bookkeeping (product code: no oracle import).

  grid_2d        cpp/tests/utilities/mg_utilities.cpp:58-61, python dask/comms/comms.py:40-44
  partition_t    cpp/include/cugraph/graph_view.hpp:81-242
"""
from __future__ import annotations


def grid_2d(comm_size: int) -> tuple[int, int]:
    """(row_comm_size R, col_comm_size C): R = largest divisor of P not above sqrt(P).
    P=2 -> 1x2, 4 -> 2x2, 8 -> 2x4."""
    r = int(comm_size ** 0.5)
    while comm_size % r != 0:
        r -= 1
    return r, comm_size // r


def rank_coords(rank: int, row_comm_size: int) -> tuple[int, int]:
    """(row_comm_rank, col_comm_rank) of a global rank (partition_manager.hpp:86-103)."""
    return rank % row_comm_size, rank // row_comm_size


def row_comm_ranks(rank: int, comm_size: int) -> list[int]:
    """world ranks of `rank`'s row communicator, in row_comm rank order: the ranks sharing
    rank // R (partition_manager.hpp:92-96).  The gathered vector is all-gathered over these."""
    r, _ = grid_2d(comm_size)
    base = (rank // r) * r
    return [base + j for j in range(r)]


def col_comm_ranks(rank: int, comm_size: int) -> list[int]:
    """world ranks of `rank`'s col communicator, in col_comm rank order: the ranks sharing
    rank % R (partition_manager.hpp:97-100).  Partial sums are reduced over these."""
    r, c = grid_2d(comm_size)
    return [i * r + rank % r for i in range(c)]


class Partition:
    """partition_t (graph_view.hpp:81-242): the ranges one rank works on, from the vertex
    partition offsets of all ranks."""

    def __init__(self, vertex_partition_offsets, comm_size: int, rank: int):
        self.offsets = [int(x) for x in vertex_partition_offsets]
        assert len(self.offsets) == comm_size + 1
        self.comm_size, self.rank = comm_size, rank
        self.row_comm_size, self.col_comm_size = grid_2d(comm_size)
        self.row_comm_rank, self.col_comm_rank = rank_coords(rank, self.row_comm_size)

    def vertex_partition_range(self, idx: int) -> tuple[int, int]:
        return self.offsets[idx], self.offsets[idx + 1]

    def local_vertex_partition_range(self) -> tuple[int, int]:
        return self.vertex_partition_range(self.rank)

    def number_of_local_edge_partitions(self) -> int:
        return self.col_comm_size

    def local_edge_partition_major_range(self, i: int) -> tuple[int, int]:
        """majors of local edge partition i = vertex partition R * i + row_comm_rank (:181-189)"""
        return self.vertex_partition_range(self.row_comm_size * i + self.row_comm_rank)

    def local_edge_partition_minor_range(self) -> tuple[int, int]:
        """minors = the R vertex partitions of this rank's grid column (:204-212)"""
        r = self.row_comm_size
        return self.offsets[self.col_comm_rank * r], self.offsets[(self.col_comm_rank + 1) * r]

    def major_owner(self, i: int) -> int:
        """world rank that owns the majors of local edge partition i (= col_comm rank i)"""
        return self.row_comm_size * i + self.row_comm_rank
