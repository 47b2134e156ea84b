"""debug: tiny graph on N ranks, print what the library built next to the oracle"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests import mg_worker, golden_vectors as gv
from cugraph_forked_b200 import engine
from oracle import graph_oracle as go
handle, rank, world = mg_worker.make_handle()
src, dst, wgt = gv.G6_SRC, gv.G6_DST, gv.G6_WGT
mode = os.environ.get("MODE", "rank0")
if mode == "rank0":
    sel = slice(0, len(src)) if rank == 0 else slice(0, 0)
else:
    sel = slice(rank, len(src), world)
t = lambda a: torch.as_tensor(a[sel]).cuda()
g = engine.Graph(handle, t(src), t(dst), t(wgt), multi_gpu=True)
o = go.build_mg(src, dst, wgt, world)
print(rank, "lib vpo", g.vertex_partition_offsets(), "oracle", list(o["vertex_partition_offsets"]),
      "V", g.num_vertices, "labels", g.renumber_map().cpu().numpy(), "oracle", o["labels"][rank], flush=True)
