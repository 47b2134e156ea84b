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


def _emulated(path, csv=False, delimiter=" ", skip_rows=0, vertex_dtype=torch.int32,
              weight_dtype=torch.float32, slow_capacity=-1):
    """The device parser's per-thread bodies (csrc/ingest_gpu.cuh) run tile by tile on the host
    through the library's test hook b200_debug_ingest_emulate."""
    import ctypes
    fn = _lib.lib().b200_debug_ingest_emulate
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_char, ctypes.c_int, ctypes.c_int,
                   ctypes.c_int, ctypes.c_longlong, ctypes.POINTER(ctypes.c_void_p)]
    vt, wt = io._types(vertex_dtype, weight_dtype)
    el = ctypes.c_void_p()
    _lib.check(fn(os.fsencode(path), int(csv), delimiter.encode(), skip_rows, vt, wt, slow_capacity,
                  ctypes.byref(el)))
    return io._finish(el, "cpu")


class _Reader:
    """read_matrix_market / read_csv with the host parser or the emulated device parser."""

    def __init__(self, kind):
        self.kind = kind

    def mtx(self, path, threads=0, **kw):
        if self.kind == "host":
            return io.read_matrix_market(path, device="cpu", threads=threads, **kw)
        return _emulated(path, **kw)

    def csv(self, path, delimiter=" ", skip_rows=0, threads=0, **kw):
        if self.kind == "host":
            return io.read_csv(path, delimiter=delimiter, skip_rows=skip_rows, device="cpu",
                               threads=threads, **kw)
        return _emulated(path, csv=True, delimiter=delimiter, skip_rows=skip_rows, **kw)


@pytest.fixture(params=["host", "device-emulated"])
def reader(request):
    return _Reader("host" if request.param == "host" else "emulated")


def _write(tmp_path, name, text):
    p = tmp_path / name
    p.write_bytes(text.encode() if isinstance(text, str) else text)
    return str(p)


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("wdt", [torch.float32, torch.float64])
@pytest.mark.parametrize("threads", [1, 3, 16])
def test_matrix_market_matches_reference_reader(reader, name, wdt, threads):
    path = os.path.join(GOLDEN, name)
    npdt = np.float32 if wdt == torch.float32 else np.float64
    rs, rd, rw, nv, sym = graph_oracle.read_mtx(path, npdt)
    src, dst, w, info = reader.mtx(path, weight_dtype=wdt, threads=threads)
    assert src.dtype == torch.int32 and w.dtype == wdt
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw)  # bit-exact: same double -> weight_t narrowing
    assert info.num_vertices == nv and info.is_symmetric == sym
    assert info.num_edges == len(rs)


@pytest.mark.parametrize("name", ["karate.mtx", "dolphins.mtx", "netscience.mtx"])
@pytest.mark.parametrize("wdt", [torch.float32, torch.float64])
def test_matrix_market_matches_the_compiled_reference(oracle_built, reader, name, wdt):
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
    src, dst, w, info = reader.mtx(path, weight_dtype=wdt, threads=4)
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw)
    assert (info.num_vertices, info.is_symmetric, info.is_pattern) == (nv, sym, pat)


def test_generated_files_match_the_compiled_reference(oracle_built, reader, tmp_path):
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
        src, dst, w, info = reader.mtx(p, threads=6)
        assert info.parser == ("host" if reader.kind == "host" else "device")
        assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
        assert np.array_equal(w.numpy(), rw)
        assert (info.num_vertices, info.is_symmetric, info.is_pattern) == (nv, sym, pat)


def test_matrix_market_format_cases(reader, tmp_path):
    # general, pattern, comments and blank lines anywhere, CRLF, '+' signs, tabs
    p = _write(tmp_path, "a.mtx",
               "%%MatrixMarket matrix coordinate pattern general\r\n% a comment\r\n\r\n"
               "4 4 3\r\n1 2\r\n% inside\r\n\r\n +3\t4 \r\n4 4\r\n")
    src, dst, w, info = reader.mtx(p)
    assert src.tolist() == [0, 2, 3] and dst.tolist() == [1, 3, 3] and w.tolist() == [1.0] * 3
    assert info.is_pattern and not info.is_symmetric and info.file_entries == 3
    # symmetric integer values, diagonal entries are not mirrored, no trailing newline
    p = _write(tmp_path, "b.mtx",
               "%%MatrixMarket matrix coordinate integer symmetric\n3 3 3\n2 1 5\n2 2 7\n3 1 -2")
    src, dst, w, info = reader.mtx(p, vertex_dtype=torch.int64, weight_dtype=torch.float64)
    assert src.dtype == torch.int64
    assert src.tolist() == [1, 0, 1, 2, 0] and dst.tolist() == [0, 1, 1, 0, 2]
    assert w.tolist() == [5.0, 5.0, 7.0, -2.0, -2.0]
    assert info.is_symmetric and info.num_edges == 5 and info.file_entries == 3
    # real values with exponents; weight_dtype=None drops them (test_weighted = false)
    p = _write(tmp_path, "c.mtx",
               "%%MatrixMarket matrix coordinate real general\n2 2 2\n1 2 1.5e-3\n2 1 .25\n")
    _, _, w, _ = reader.mtx(p)
    assert w.tolist() == [np.float32(1.5e-3), 0.25]
    src, dst, w, info = reader.mtx(p, weight_dtype=None)
    assert w is None and src.tolist() == [0, 1]
    # an empty matrix
    p = _write(tmp_path, "d.mtx", "%%MatrixMarket matrix coordinate real general\n5 5 0\n")
    src, dst, w, info = reader.mtx(p)
    assert src.numel() == 0 and w.numel() == 0 and info.num_vertices == 5


def test_matrix_market_errors(reader, tmp_path):
    with pytest.raises(_lib.CugraphError, match="fopen .* failure"):
        reader.mtx(str(tmp_path / "missing.mtx"))
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
            reader.mtx(_write(tmp_path, name + ".mtx", text))
    big = "%%MatrixMarket matrix coordinate pattern general\n3000000000 3000000000 1\n1 2999999999\n"
    p = _write(tmp_path, "big.mtx", big)
    with pytest.raises(_lib.CugraphError, match="does not fit vertex_type INT32"):
        reader.mtx(p)
    src, dst, _, _ = reader.mtx(p, vertex_dtype=torch.int64)
    assert dst.tolist() == [2999999998]
    with pytest.raises(TypeError):
        io.read_matrix_market(p, vertex_dtype=torch.float32, device="cpu")


@pytest.mark.parametrize("delimiter", [" ", ",", "\t"])
def test_csv_matches_pandas(reader, tmp_path, delimiter):
    rng = np.random.default_rng(1)
    n = 50000
    df = pd.DataFrame({"s": rng.integers(0, 100000, n), "d": rng.integers(0, 100000, n),
                       "w": rng.random(n).astype(np.float32)})
    p = str(tmp_path / "e.csv")
    df.to_csv(p, sep=delimiter, header=False, index=False)
    want = pd.read_csv(p, delimiter=delimiter, header=None,
                       dtype={0: "int32", 1: "int32", 2: "float32"})
    one = reader.csv(p, delimiter=delimiter, threads=1)
    many = reader.csv(p, delimiter=delimiter, threads=13)
    for src, dst, w, info in (one, many):
        assert np.array_equal(src.numpy(), want[0].to_numpy())
        assert np.array_equal(dst.numpy(), want[1].to_numpy())
        assert np.array_equal(w.numpy(), want[2].to_numpy())
        assert info.num_edges == n and info.num_vertices == int(max(want[0].max(), want[1].max())) + 1
    # header row skipped; two-column read ignores the weights
    df.to_csv(p, sep=delimiter, header=True, index=False)
    src, dst, w, _ = reader.csv(p, delimiter=delimiter, skip_rows=1, weight_dtype=None)
    assert w is None and np.array_equal(src.numpy(), want[0].to_numpy())
    with pytest.raises(_lib.CugraphError, match=r"could not read matrix data \(line 1\)"):
        reader.csv(p, delimiter=delimiter)  # the header is not an edge


def test_csv_karate_dataset_shape(reader, tmp_path):
    """The reference's datasets are space-separated `src dst weight` files read with
    dtype=['int32','int32','float32'] (datasets/karate.csv); build one from the golden MTX."""
    rs, rd, rw, nv, _ = graph_oracle.read_mtx(os.path.join(GOLDEN, "karate.mtx"), np.float32)
    text = "".join(f"{s} {d} {w:.1f}\n" for s, d, w in zip(rs, rd, rw))
    p = _write(tmp_path, "karate.csv", text)
    src, dst, w, info = reader.csv(p)
    assert np.array_equal(src.numpy(), rs) and np.array_equal(dst.numpy(), rd)
    assert np.array_equal(w.numpy(), rw) and info.num_vertices == nv
    with pytest.raises(_lib.CugraphError, match="could not read matrix data"):
        reader.csv(_write(tmp_path, "two.csv", "0 1\n1 2\n"))  # weight expected
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


# ---- the device parser's own corners (emulated on the host) ---------------------------------

def test_device_parser_values_are_correctly_rounded(tmp_path):
    """Decimal strings of every shape: short ones take the exact path on the device, the rest are
    flagged and re-parsed on the host; either way the double is the one float() gives."""
    rng = np.random.default_rng(9)
    texts = ["0", "-0.0", "+3", "5.", ".5", "1e3", "1E+3", "1.e5", "2.5e-3", "123456789012345",
             "1234567890123456", "12345678901234567890", "0.000000000000000000001", "1e22", "1e23",
             "1e-22", "1e-23", "9007199254740993", "0.1", "0.30000000000000004", "1e", "7e+",
             "000123.4500", "1.7976931348623157e308", "4.9e-324", "00000000000000000000001.5"]
    for _ in range(3000):
        digits = int(rng.integers(1, 20))
        m = "".join(rng.choice(list("0123456789"), digits))
        cut = int(rng.integers(0, digits + 1))
        t = (m[:cut] or "0") + "." + m[cut:] if rng.random() < 0.8 else m
        if rng.random() < 0.4:
            t += "e" + str(int(rng.integers(-30, 31)))
        if rng.random() < 0.3:
            t = "-" + t
        texts.append(t)
    expect = []
    for t in texts:
        core = t.rstrip("+").rstrip("eE") if t.endswith(("e", "E", "e+", "E+")) else t
        expect.append(float(core))
    p = _write(tmp_path, "v.csv", "".join(f"{i} {i + 1} {t}\n" for i, t in enumerate(texts)))
    for cap in (-1, 7):  # 7: more flagged values than the patch list holds -> host reader
        _, _, w, info = _emulated(p, csv=True, weight_dtype=torch.float64, slow_capacity=cap)
        assert np.array_equal(w.numpy(), np.array(expect))
        assert info.parser == ("device" if cap < 0 else "host")
    _, _, w32, _ = _emulated(p, csv=True)
    assert np.array_equal(w32.numpy(), np.array(expect).astype(np.float32))
    host = io.read_csv(p, weight_dtype=torch.float64, device="cpu")
    assert np.array_equal(host[2].numpy(), np.array(expect))


def test_device_parser_tile_and_chunk_boundaries(tmp_path):
    """Lines of every length up to the look-ahead limit, so that line starts, line ends and
    values fall on chunk (260 B) and tile (65 KB) boundaries; one longer line sends the file to
    the host reader."""
    rng = np.random.default_rng(4)
    lines, want = [], []
    for i in range(6000):
        r, c = int(rng.integers(1, 10**6)), int(rng.integers(1, 10**6))
        v = round(float(rng.random()), 6)
        pad = " " * int(rng.integers(0, 5) if i % 50 else rng.integers(200, 490))
        lines.append(f"{pad}{r}{pad[:7]} {c} {v}{pad}\n")
        want.append((r - 1, c - 1, v))
        if i % 97 == 0:
            lines.append("%" + "c" * int(rng.integers(0, 900)) + "\n")
        if i % 131 == 0:
            lines.append("\n")
    head = "%%MatrixMarket matrix coordinate real general\n1000000 1000000 6000\n"
    p = _write(tmp_path, "b.mtx", head + "".join(lines))
    assert os.path.getsize(p) > 3 * 65536
    host = io.read_matrix_market(p, weight_dtype=torch.float64, device="cpu")
    emu = _emulated(p, weight_dtype=torch.float64)
    assert emu[3].parser == "device" and host[3].parser == "host"
    for a, b in zip(host[:3], emu[:3]):
        assert torch.equal(a, b)
    assert emu[0].tolist() == [r for r, _, _ in want] and emu[2].tolist() == [v for _, _, v in want]
    # a 5000-byte comment line is longer than the kernels look ahead
    lines.insert(3000, "%" + "x" * 5000 + "\n")
    p = _write(tmp_path, "c.mtx", head + "".join(lines))
    emu = _emulated(p, weight_dtype=torch.float64)
    assert emu[3].parser == "host"
    for a, b in zip(host[:3], emu[:3]):
        assert torch.equal(a, b)


def test_device_parser_large_symmetric_file(tmp_path):
    rng = np.random.default_rng(6)
    n = 150000
    r, c = rng.integers(1, 5001, n), rng.integers(1, 5001, n)
    v = rng.standard_normal(n)
    body = "".join(f"{a} {b} {x:.9g}\n" for a, b, x in zip(r, c, v))
    p = _write(tmp_path, "s.mtx", f"%%MatrixMarket matrix coordinate real symmetric\n5000 5000 {n}\n" + body)
    host = io.read_matrix_market(p, weight_dtype=torch.float64, device="cpu", threads=3)
    emu = _emulated(p, weight_dtype=torch.float64)
    for a, b in zip(host[:3], emu[:3]):
        assert torch.equal(a, b)
    assert emu[3].num_edges == n + int((r != c).sum()) and emu[3].is_symmetric
    assert emu[3].parser == "device"
