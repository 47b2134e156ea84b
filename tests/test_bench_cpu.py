"""CPU: the reference arm of bench.py (`--impl reference`) and the fast input builders it uses.
The arm must print one JSON line with the contract's keys, run the SAME workload string as the
GPU arm, and finish quickly at a small scale."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line(oracle_built):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                        "--scale", "12", "--steps", "4", "--warmup", "3"],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step",
              "higher_is_better", "scaling", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["impl"] == "reference" and line["unit"] == "GTEPS" and line["value"] > 0
    assert line["config"]["same_config"] is True
    assert line["cpu_baseline"]["iterations"] == 4 and line["cpu_baseline"]["cores"] == 1
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    sys.path.insert(0, ROOT)
    import bench
    assert line["config"]["workload"] == bench.workload_string(12, "f32")


def test_bounded_iterations():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.bounded_iterations(24, 20) == 20          # 20 iterations of scale 24 fit 2 minutes
    assert 1 <= bench.bounded_iterations(28, 20) < 20      # scale 28 does not
    assert bench.bounded_iterations(20, 3, 40.0) == 3


def test_fast_builders_equal_the_numpy_ones(oracle_built):
    from oracle import cpu_oracle, graph_oracle
    scale, ne = 11, 16 << 11
    a = graph_oracle.rmat_edgelist(scale, ne, seed=3, first_edge=17)
    b = cpu_oracle.rmat_edgelist_fast(scale, ne, seed=3, first_edge=17, threads=3)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    o, i, _ = cpu_oracle.csc_from_coo(a[0], a[1], None, 1 << scale)
    o2, i2 = cpu_oracle.csc_from_coo_fast(a[0], a[1], 1 << scale, threads=3)
    assert np.array_equal(o, o2) and np.array_equal(i, i2)
