"""Host-side timing of the Matrix-Market reader (csrc/ingest.cu) beside the reference's own
reader (oracle/_ref: mm_properties + mm_to_coo, fscanf per entry).  No GPU involved.
usage: python scripts/ingest_bench.py [entries] [threads...]"""
import os
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cugraph_forked_b200 import io  # noqa: E402
from oracle import cpu_oracle  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
threads = [int(t) for t in sys.argv[2:]] or [1, 4, 0]
rng = np.random.default_rng(0)
nv = 1 << 20
r, c = rng.integers(1, nv + 1, n), rng.integers(1, nv + 1, n)
v = rng.random(n)
with tempfile.TemporaryDirectory(dir=os.environ.get("INGEST_TMP")) as d:
    path = os.path.join(d, "g.mtx")
    with open(path, "w") as f:
        f.write(f"%%MatrixMarket matrix coordinate real general\n{nv} {nv} {n}\n")
        np.savetxt(f, np.column_stack([r, c, v]), fmt="%d %d %.6f")
    size = os.path.getsize(path)
    print(f"{n} entries, {size / 1e6:.1f} MB, host threads available: {os.cpu_count()}")
    if cpu_oracle.ref_available():
        t0 = time.perf_counter()
        ref = cpu_oracle.read_mtx_ref(path, np.float32)
        dt = time.perf_counter() - t0
        print(f"reference reader (1 thread): {dt:.3f} s  {n / dt / 1e6:.2f} M entries/s  {size / dt / 1e6:.0f} MB/s")
    for t in threads:
        io.read_matrix_market(path, device="cpu", threads=t)  # page cache + first-touch warm-up
        t0 = time.perf_counter()
        got = io.read_matrix_market(path, device="cpu", threads=t)
        dt = time.perf_counter() - t0
        ok = (not cpu_oracle.ref_available()) or (
            np.array_equal(got[0].numpy(), ref[0]) and np.array_equal(got[2].numpy(), ref[2]))
        print(f"library reader, threads={t or 'all'}: {dt:.3f} s  {n / dt / 1e6:.2f} M entries/s  "
              f"{size / dt / 1e6:.0f} MB/s  identical={ok}")
    if torch.cuda.is_available():
        # host reader + upload vs the device parser (B200_INGEST_VERBOSE=1 prints its phases)
        from cugraph_forked_b200 import engine
        h = engine.Handle()
        for parser in ("host", "device"):
            io.read_matrix_market(path, parser=parser, handle=h) if parser == "device" else \
                io.read_matrix_market(path)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            got = io.read_matrix_market(path, parser=parser, handle=h) if parser == "device" else \
                io.read_matrix_market(path)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            ok = (not cpu_oracle.ref_available()) or (
                np.array_equal(got[0].cpu().numpy(), ref[0]) and
                np.array_equal(got[2].cpu().numpy(), ref[2]))
            print(f"file -> device columns, parser={parser} ({got[3].parser}): {dt:.3f} s  "
                  f"{n / dt / 1e6:.2f} M entries/s  {size / dt / 1e6:.0f} MB/s  identical={ok}")
