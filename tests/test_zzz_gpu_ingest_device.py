"""GPU: the device-side text parser (csrc/ingest_gpu.cu, io.read_*(parser="device")) against the
host reader, which tests/test_ingest.py holds bit-identical to the reference's own reader.  The
per-thread parsing code is also run on the CPU there (emulated grid); this file covers what only
a GPU can: the kernels' staging, scans and launches.  (Named to run after everything else: these
kernels were written after round 1's GPU budget was spent.)"""
import os

import numpy as np
import pytest
import torch

# First hardware run: round 2, call r2a (profiles/r2_notes.md) - 11 cases passed; the opt-in
# guard of round 1 is gone.
pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _same(a, b):
    for x, y in zip(a[:3], b[:3]):
        if x is None or y is None:
            assert x is None and y is None
        else:
            assert x.dtype == y.dtype and torch.equal(x.cpu(), y.cpu())
    for k in ("num_rows", "num_cols", "num_edges", "file_entries", "is_symmetric", "is_pattern"):
        assert getattr(a[3], k) == getattr(b[3], k), k


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("wdt", [torch.float32, torch.float64, None])
def test_golden_files(handle, name, wdt):
    from cugraph_forked_b200 import io
    path = os.path.join(GOLDEN, name)
    dev = io.read_matrix_market(path, weight_dtype=wdt, parser="device", handle=handle)
    assert dev[0].is_cuda and dev[3].parser == "device"
    _same(dev, io.read_matrix_market(path, weight_dtype=wdt, device="cpu"))


def test_generated_files(handle, tmp_path):
    from cugraph_forked_b200 import io
    rng = np.random.default_rng(8)
    n = 400000
    r, c = rng.integers(1, 200001, n), rng.integers(1, 200001, n)
    v = rng.standard_normal(n)
    body = "".join(f"{a} {b} {x:.9g}\n" for a, b, x in zip(r, c, v))
    for symmetry in ("general", "symmetric"):
        p = str(tmp_path / f"{symmetry}.mtx")
        with open(p, "w") as f:
            f.write(f"%%MatrixMarket matrix coordinate real {symmetry}\n% generated\n200000 200000 {n}\n")
            f.write(body)
        for vdt in (torch.int32, torch.int64):
            dev = io.read_matrix_market(p, vertex_dtype=vdt, weight_dtype=torch.float64,
                                        parser="device", handle=handle)
            assert dev[3].parser == "device"
            _same(dev, io.read_matrix_market(p, vertex_dtype=vdt, weight_dtype=torch.float64,
                                             device="cpu"))
    # values with 17 significant digits are re-parsed on the host and patched in
    p = str(tmp_path / "long.csv")
    with open(p, "w") as f:
        f.write("".join(f"{a},{b},{x:.17g}\n" for a, b, x in zip(r[:50000], c[:50000], v[:50000])))
    dev = io.read_csv(p, delimiter=",", weight_dtype=torch.float64, parser="device", handle=handle)
    assert dev[3].parser == "device" and np.array_equal(dev[2].cpu().numpy(), v[:50000])
    _same(dev, io.read_csv(p, delimiter=",", weight_dtype=torch.float64, device="cpu"))


def test_errors_and_host_hand_over(handle, tmp_path):
    from cugraph_forked_b200 import _lib, io
    p = str(tmp_path / "bad.mtx")
    with open(p, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n3 3 2\n1 2 1\nx y z\n")
    with pytest.raises(_lib.CugraphError, match=r"could not read matrix data \(line 4\)"):
        io.read_matrix_market(p, parser="device", handle=handle)
    p = str(tmp_path / "long.mtx")
    with open(p, "w") as f:
        f.write("%%MatrixMarket matrix coordinate pattern general\n3 3 2\n1 2\n%" + "x" * 5000 + "\n2 3\n")
    dev = io.read_matrix_market(p, parser="device", handle=handle)
    assert dev[3].parser == "host" and dev[0].is_cuda and dev[0].tolist() == [0, 1]
    p = str(tmp_path / "empty.mtx")
    with open(p, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n5 5 0\n")
    dev = io.read_matrix_market(p, parser="device", handle=handle)
    assert dev[0].numel() == 0 and dev[3].num_vertices == 5
