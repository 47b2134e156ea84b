"""CPU: the library's edge-list readers (csrc/ingest.cu through cugraph_forked_b200/io.py; host
code, no device needed) against the oracle's restatement of the reference reader
(oracle/graph_oracle.py read_mtx <- matrix_market_file_utilities.cu mm_to_coo) on the reference's
own datasets, against pandas.read_csv for delimited files, and on the edge cases of the format."""
import os

import numpy as np
import pandas as pd
import pytest
import torch

from cugraph_forked_b200 import _lib, io
from oracle import graph_oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _write(tmp_path, name, text):
    p = tmp_path / name
    p.write_bytes(text.encode() if isinstance(text, str) else text)
    return str(p)


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("wdt", [torch.float32, torch.float64])
@pytest.mark.parametrize("threads", [1, 3, 16])
def test_matrix_market_matches_reference_reader(name, wdt, threads):
    path = os.path.join(GOLDEN, name)
    npdt = np.float32 if wdt == torch.float32 else np.float64
    rs, rd, rw, nv, sym = graph_oracle.read_mtx(path, npdt)
    src, dst, w, info = io.read_matrix_market(path, weight_dtype=wdt, device="cpu", threads=threads)
    assert src.dtype == torch.int32 and w.dtype == wdt
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw)  # bit-exact: same double -> weight_t narrowing
    assert info.num_vertices == nv and info.is_symmetric == sym
    assert info.num_edges == len(rs)


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("wdt", [torch.float32, torch.float64])
def test_matrix_market_matches_the_compiled_reference(oracle_built, name, wdt):
    """oracle/_ref holds the reference's own mm_properties / mm_to_coo: pins both the oracle's
    restatement and the library's reader."""
    if not oracle_built.ref_available():
        pytest.skip("oracle/_ref not built")
    path = os.path.join(GOLDEN, name)
    npdt = np.float32 if wdt == torch.float32 else np.float64
    rs, rd, rw, nv, sym, pat = oracle_built.read_mtx_ref(path, npdt)
    os_, od, ow, onv, osym = graph_oracle.read_mtx(path, npdt)
    assert np.array_equal(os_, rs) and np.array_equal(od, rd) and np.array_equal(ow, rw)
    assert onv == nv and osym == sym
    src, dst, w, info = io.read_matrix_market(path, weight_dtype=wdt, device="cpu", threads=4)
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw)
    assert (info.num_vertices, info.is_symmetric, info.is_pattern) == (nv, sym, pat)


def test_generated_files_match_the_compiled_reference(oracle_built, tmp_path):
    if not oracle_built.ref_available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(5)
    for field, symmetry in (("real", "general"), ("real", "symmetric"), ("integer", "symmetric"),
                            ("pattern", "general"), ("pattern", "symmetric")):
        n = 20000
        r, c = rng.integers(1, 3001, n), rng.integers(1, 3001, n)
        if field == "real":
            body = "".join(f"{a} {b} {x:.9g}\n" for a, b, x in zip(r, c, rng.standard_normal(n)))
        elif field == "integer":
            body = "".join(f"{a} {b} {x}\n" for a, b, x in zip(r, c, rng.integers(-50, 50, n)))
        else:
            body = "".join(f"{a} {b}\n" for a, b in zip(r, c))
        p = _write(tmp_path, f"{field}_{symmetry}.mtx",
                   f"%%MatrixMarket matrix coordinate {field} {symmetry}\n%generated\n3000 3000 {n}\n"
                   + body)
        rs, rd, rw, nv, sym, pat = oracle_built.read_mtx_ref(p, np.float32)
        src, dst, w, info = io.read_matrix_market(p, device="cpu", threads=6)
        assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
        assert np.array_equal(w.numpy(), rw)
        assert (info.num_vertices, info.is_symmetric, info.is_pattern) == (nv, sym, pat)


def test_matrix_market_format_cases(tmp_path):
    # general, pattern, comments and blank lines anywhere, CRLF, '+' signs, tabs
    p = _write(tmp_path, "a.mtx",
               "%%MatrixMarket matrix coordinate pattern general\r\n% a comment\r\n\r\n"
               "4 4 3\r\n1 2\r\n% inside\r\n\r\n +3\t4 \r\n4 4\r\n")
    src, dst, w, info = io.read_matrix_market(p, device="cpu")
    assert src.tolist() == [0, 2, 3] and dst.tolist() == [1, 3, 3] and w.tolist() == [1.0] * 3
    assert info.is_pattern and not info.is_symmetric and info.file_entries == 3
    # symmetric integer values, diagonal entries are not mirrored, no trailing newline
    p = _write(tmp_path, "b.mtx",
               "%%MatrixMarket matrix coordinate integer symmetric\n3 3 3\n2 1 5\n2 2 7\n3 1 -2")
    src, dst, w, info = io.read_matrix_market(p, vertex_dtype=torch.int64,
                                              weight_dtype=torch.float64, device="cpu")
    assert src.dtype == torch.int64
    assert src.tolist() == [1, 0, 1, 2, 0] and dst.tolist() == [0, 1, 1, 0, 2]
    assert w.tolist() == [5.0, 5.0, 7.0, -2.0, -2.0]
    assert info.is_symmetric and info.num_edges == 5 and info.file_entries == 3
    # real values with exponents; weight_dtype=None drops them (test_weighted = false)
    p = _write(tmp_path, "c.mtx",
               "%%MatrixMarket matrix coordinate real general\n2 2 2\n1 2 1.5e-3\n2 1 .25\n")
    _, _, w, _ = io.read_matrix_market(p, device="cpu")
    assert w.tolist() == [np.float32(1.5e-3), 0.25]
    src, dst, w, info = io.read_matrix_market(p, weight_dtype=None, device="cpu")
    assert w is None and src.tolist() == [0, 1]
    # an empty matrix
    p = _write(tmp_path, "d.mtx", "%%MatrixMarket matrix coordinate real general\n5 5 0\n")
    src, dst, w, info = io.read_matrix_market(p, device="cpu")
    assert src.numel() == 0 and w.numel() == 0 and info.num_vertices == 5


def test_matrix_market_errors(tmp_path):
    with pytest.raises(_lib.CugraphError, match="fopen .* failure"):
        io.read_matrix_market(str(tmp_path / "missing.mtx"), device="cpu")
    cases = {
        "complex": ("%%MatrixMarket matrix coordinate complex general\n1 1 1\n1 1 1 0\n",
                    "invalid Matrix Market file properties"),
        "skew": ("%%MatrixMarket matrix coordinate real skew-symmetric\n2 2 1\n2 1 1\n",
                 "invalid Matrix Market file properties"),
        "array": ("%%MatrixMarket matrix array real general\n1 1\n1\n",
                  "invalid Matrix Market file properties"),
        "banner": ("1 1 1\n1 1 1\n", "could not read Matrix Market file properties"),
        "size": ("%%MatrixMarket matrix coordinate real general\n% only comments\n",
                 "could not read Matrix Market file properties"),
        "short": ("%%MatrixMarket matrix coordinate real general\n3 3 2\n1 2 1\n",
                  "1 entries, size line says 2"),
        "garbage": ("%%MatrixMarket matrix coordinate real general\n3 3 2\n1 2 1\nx y z\n",
                    r"could not read matrix data \(line 4\)"),
        "novalue": ("%%MatrixMarket matrix coordinate real general\n3 3 1\n1 2\n",
                    r"could not read matrix data \(line 3\)"),
        "beyond": ("%%MatrixMarket matrix coordinate pattern general\n3 3 1\n1 9\n",
                   "index beyond the size line"),
        "zero": ("%%MatrixMarket matrix coordinate pattern general\n3 3 1\n0 1\n",
                 "negative vertex id"),
    }
    for name, (text, msg) in cases.items():
        with pytest.raises(_lib.CugraphError, match=msg):
            io.read_matrix_market(_write(tmp_path, name + ".mtx", text), device="cpu")
    big = "%%MatrixMarket matrix coordinate pattern general\n3000000000 3000000000 1\n1 2999999999\n"
    p = _write(tmp_path, "big.mtx", big)
    with pytest.raises(_lib.CugraphError, match="does not fit vertex_type INT32"):
        io.read_matrix_market(p, device="cpu")
    src, dst, _, _ = io.read_matrix_market(p, vertex_dtype=torch.int64, device="cpu")
    assert dst.tolist() == [2999999998]
    with pytest.raises(TypeError):
        io.read_matrix_market(p, vertex_dtype=torch.float32, device="cpu")


@pytest.mark.parametrize("delimiter", [" ", ",", "\t"])
def test_csv_matches_pandas(tmp_path, delimiter):
    rng = np.random.default_rng(1)
    n = 50000
    df = pd.DataFrame({"s": rng.integers(0, 100000, n), "d": rng.integers(0, 100000, n),
                       "w": rng.random(n).astype(np.float32)})
    p = str(tmp_path / "e.csv")
    df.to_csv(p, sep=delimiter, header=False, index=False)
    want = pd.read_csv(p, delimiter=delimiter, header=None,
                       dtype={0: "int32", 1: "int32", 2: "float32"})
    one = io.read_csv(p, delimiter=delimiter, device="cpu", threads=1)
    many = io.read_csv(p, delimiter=delimiter, device="cpu", threads=13)
    for src, dst, w, info in (one, many):
        assert np.array_equal(src.numpy(), want[0].to_numpy())
        assert np.array_equal(dst.numpy(), want[1].to_numpy())
        assert np.array_equal(w.numpy(), want[2].to_numpy())
        assert info.num_edges == n and info.num_vertices == int(max(want[0].max(), want[1].max())) + 1
    # header row skipped; two-column read ignores the weights
    df.to_csv(p, sep=delimiter, header=True, index=False)
    src, dst, w, _ = io.read_csv(p, delimiter=delimiter, skip_rows=1, weight_dtype=None,
                                 device="cpu")
    assert w is None and np.array_equal(src.numpy(), want[0].to_numpy())
    with pytest.raises(_lib.CugraphError, match=r"could not read matrix data \(line 1\)"):
        io.read_csv(p, delimiter=delimiter, device="cpu")  # the header is not an edge


def test_csv_karate_dataset_shape(tmp_path):
    """The reference's datasets are space-separated `src dst weight` files read with
    dtype=['int32','int32','float32'] (datasets/karate.csv); build one from the golden MTX."""
    rs, rd, rw, nv, _ = graph_oracle.read_mtx(os.path.join(GOLDEN, "karate.mtx"), np.float32)
    text = "".join(f"{s} {d} {w:.1f}\n" for s, d, w in zip(rs, rd, rw))
    p = _write(tmp_path, "karate.csv", text)
    src, dst, w, info = io.read_csv(p, device="cpu")
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw) and info.num_vertices == nv
    with pytest.raises(_lib.CugraphError, match="could not read matrix data"):
        io.read_csv(_write(tmp_path, "two.csv", "0 1\n1 2\n"), device="cpu")  # weight expected
    with pytest.raises(ValueError, match="one character"):
        io.read_csv(p, delimiter="::", device="cpu")
    with pytest.raises(_lib.CugraphError, match="invalid delimiter"):
        io.read_csv(p, delimiter="1", device="cpu")


def test_thread_count_does_not_change_the_result(tmp_path):
    rng = np.random.default_rng(2)
    n = 200000
    r, c = rng.integers(1, 5001, n), rng.integers(1, 5001, n)
    v = rng.standard_normal(n)
    lines = [f"{a} {b} {x:.17g}\n" for a, b, x in zip(r, c, v)]
    p = _write(tmp_path, "r.mtx",
               f"%%MatrixMarket matrix coordinate real symmetric\n5000 5000 {n}\n" + "".join(lines))
    base = io.read_matrix_market(p, weight_dtype=torch.float64, device="cpu", threads=1)
    off = r != c
    assert base[3].num_edges == n + int(off.sum())
    # file order with the mirrored entry right after each off-diagonal entry
    pos = np.cumsum(np.concatenate([[0], 1 + off[:-1]]))
    assert np.array_equal(base[0].numpy()[pos], r - 1) and np.array_equal(base[1].numpy()[pos], c - 1)
    assert np.array_equal(base[2].numpy()[pos], v)  # %.17g round-trips doubles exactly
    assert np.array_equal(base[0].numpy()[pos[off] + 1], (c - 1)[off])
    for threads in (2, 5, 0):
        got = io.read_matrix_market(p, weight_dtype=torch.float64, device="cpu", threads=threads)
        for a, b in zip(base[:3], got[:3]):
            assert torch.equal(a, b)
