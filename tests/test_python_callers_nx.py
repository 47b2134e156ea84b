"""CPU: the reference's cuGraph-vs-NetworkX Python tests (tests/nx_cases.py) with the CUDA library
replaced by the oracle-backed stand-in.  Besides the caller logic this checks the ORACLE against
an implementation that is independent of the reference's C++ (NetworkX), on the reference's own
datasets and with the reference tests' own tolerances."""
import pytest

from tests import fake_plc, nx_cases


@pytest.fixture()
def cugraph(monkeypatch, oracle_built):
    from cugraph_forked_b200 import cugraph as cg
    monkeypatch.setattr(cg, "_plc", fake_plc)
    monkeypatch.setattr(cg, "_DEVICE", "cpu")
    return cg


@pytest.mark.parametrize("name", nx_cases.DATASETS)
@pytest.mark.parametrize("personalization_perc", [0, 10, 50])
@pytest.mark.parametrize("has_guess", [0, 1])
@pytest.mark.parametrize("has_precomputed", [0, 1])
def test_pagerank(cugraph, name, personalization_perc, has_guess, has_precomputed):
    nx_cases.pagerank_case(cugraph, name, personalization_perc, has_guess, has_precomputed)


@pytest.mark.parametrize("name", nx_cases.DATASETS)
@pytest.mark.parametrize("personalization_perc", [0, 10])
def test_pagerank_nx(cugraph, name, personalization_perc):
    nx_cases.pagerank_case(cugraph, name, personalization_perc, 1, 0, via_nx=True)


@pytest.mark.parametrize("name", nx_cases.DATASETS)
def test_katz_centrality(cugraph, name):
    nx_cases.katz_case(cugraph, name)


@pytest.mark.parametrize("name", nx_cases.DATASETS)
def test_eigenvector_centrality(cugraph, name):
    nx_cases.eigenvector_case(cugraph, name)


@pytest.mark.parametrize("name", nx_cases.DATASETS)
def test_hits(cugraph, name):
    nx_cases.hits_case(cugraph, name)


def test_networkx_inputs(cugraph):
    nx_cases.nx_input_case(cugraph)
