"""GPU: the multi-GPU (2D-partitioned) path.  World size 1 runs in-process on any GPU box (every
collective degenerates but the whole MG code path - shuffle, hash ownership, label tables,
per-partition build, generic schedule, exchange loop - executes); world sizes 2 / 4 / 8 start
one process per GPU with torchrun and are skipped when the box has fewer GPUs.
Mirrors cpp/tests/link_analysis/mg_pagerank_test.cpp and cpp/tests/c_api/mg_pagerank_test.c."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mg1(oracle_built):
    from tests import mg_worker
    os.environ.pop("WORLD_SIZE", None)
    os.environ.pop("RANK", None)
    handle, rank, world = mg_worker.make_handle()
    assert (rank, world) == (0, 1)
    return handle


@pytest.mark.parametrize("dtype,weighted", [(np.float32, False), (np.float64, True)])
def test_mg_world_size_1(mg1, dtype, weighted):
    from tests import mg_worker
    mg_worker.run(11, 16, dtype, weighted, mg1, 0, 1, verbose=False)


@pytest.mark.parametrize("dtype,weighted", [(np.float32, True), (np.float64, False)])
def test_mg_world_size_1_banded(mg1, monkeypatch, capfd, dtype, weighted):
    """a partition large enough for the banded pre-reduction (banded.cu) on the MG code path:
    generic schedule of the reduced partition, raw epilogue, exchange loop"""
    from tests import mg_worker
    monkeypatch.setenv("B200_BANDED", "1")
    monkeypatch.setenv("B200_BAND_HOT_MIN", "1024")
    monkeypatch.setenv("B200_BAND_VERBOSE", "1")
    mg_worker.run(15, 16, dtype, weighted, mg1, 0, 1, verbose=False)
    assert "sell layout" in capfd.readouterr().err


def test_mg_tiny_graph_world_size_1(mg1):
    from tests import mg_worker
    mg_worker.run_tiny(mg1, 0, 1)


def test_mg_c_api_world_size_1(mg1):
    """cugraph_mg_graph_create + cugraph_pagerank through the pylibcugraph mirror on an MG
    handle (c_api/mg_pagerank_test.c:213-247 golden vector)."""
    from cugraph_forked_b200 import pylibcugraph as plc
    from tests import golden_vectors as gv
    rh = plc.ResourceHandle(mg1)
    t = lambda a: torch.as_tensor(a).cuda()
    g = plc.MGGraph(rh, plc.GraphProperties(), t(gv.G6_SRC), t(gv.G6_DST), t(gv.G6_WGT),
                    store_transposed=True, num_edges=len(gv.G6_SRC))
    p = gv.G6_PAGERANK
    vertices, values = plc.pagerank(rh, g, None, None, None, None, p["alpha"], p["epsilon"],
                                    p["max_iterations"], False)
    out = np.zeros(6)
    out[vertices.cpu().numpy()] = values.cpu().numpy()
    assert gv.nearly_equal(out, p["result"])
    k = gv.G6_KATZ
    vertices, values = plc.katz_centrality(rh, g, None, k["alpha"], k["beta"], k["epsilon"],
                                           k["max_iterations"], False)
    out[vertices.cpu().numpy()] = values.cpu().numpy()
    assert gv.nearly_equal(out, k["result"])


@pytest.mark.parametrize("world", [2, 4, 8])
def test_mg_torchrun(world, oracle_built):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port",
           str(29500 + world), os.path.join(ROOT, "tests", "mg_worker.py"), "--scale", "12"]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    ok = r.stdout.count("MG_WORKER_OK")
    assert r.returncode == 0 and ok == 6 * world, (r.stdout[-3000:], r.stderr[-3000:])
    # again with the banded pre-reduction forced onto the small partitions
    env = dict(os.environ, B200_BANDED="1", B200_BAND_MIN_EDGES="512", B200_BAND_HOT_MIN="64",
               B200_BAND_VERBOSE="1")
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900, env=env)
    ok = r.stdout.count("MG_WORKER_OK")
    assert r.returncode == 0 and ok == 6 * world, (r.stdout[-3000:], r.stderr[-3000:])
    assert "sell layout" in r.stderr
