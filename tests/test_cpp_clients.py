"""The C ABI from compiled clients: a pure-C program against include/cugraph_c (shaped like
cpp/tests/c_api/pagerank_test.c) and a C++ program against the cugraph::pagerank /
katz_centrality / hits shims of include/cugraph_b200/algorithms.hpp (the hits client reads its
graph through b200_read_matrix_market).  Compiling and linking needs no
GPU (CPU suite); running them is a GPU test."""
import os
import subprocess

import pytest

from cugraph_forked_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.dirname(_lib.LIB_PATH)
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")
OUT = os.path.join(ROOT, "tests", "cpp", "build")


def _build(name):
    os.makedirs(OUT, exist_ok=True)
    exe = os.path.join(OUT, name)
    src = os.path.join(ROOT, "tests", "cpp", name + (".c" if name.startswith("c_api") else ".cpp"))
    common = ["-I", os.path.join(ROOT, "include"), src, "-L", LIBDIR, "-lcugraph_b200",
              f"-Wl,-rpath,{LIBDIR}", "-o", exe]
    if src.endswith(".c"):
        cmd = ["gcc", "-std=c99", "-Wall", "-Werror"] + common + ["-lm"]
    else:
        cmd = ["g++", "-std=c++17", "-Wall", "-I", os.path.join(CUDA, "include")] + common + \
              ["-L", os.path.join(CUDA, "lib64"), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


@pytest.mark.parametrize("name", ["c_api_golden", "shim_golden", "hits_file_golden"])
def test_clients_compile_and_link(name):
    assert os.path.exists(_build(name))


@pytest.mark.gpu
@pytest.mark.parametrize("name,token", [("c_api_golden", "C_API_GOLDEN_OK"),
                                        ("shim_golden", "SHIM_GOLDEN_OK")])
def test_clients_run(name, token):
    r = subprocess.run([_build(name)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and token in r.stdout, (r.stdout, r.stderr)
# (hits_file_golden runs from tests/test_zz_gpu_python_callers.py, after the kernel parity tests)
