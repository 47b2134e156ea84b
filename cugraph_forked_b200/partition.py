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
