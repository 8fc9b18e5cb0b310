"""Pins the CPU oracle to the reference: structure arrays bit-exact, stage values, chi2 per LM
iteration (1e-9 relative) and final estimates against tests/golden/ref_*.{json,npz}, which were
produced by the unmodified reference library on a B200 (tests/golden/README.md)."""
import pytest

import golden_util
import oracle as orc


def make_oracle(prob, rk, delta):
    return orc.Oracle(prob, robust=rk, delta=delta, threads=4)


@pytest.mark.parametrize("case", golden_util.CASES)
def test_oracle_matches_reference(case):
    rep = golden_util.check_all(case, make_oracle)
    print(case, rep)
    # final estimates after 10 iterations: the oracle and the reference differ only by the
    # elimination order of the Cholesky and by atomic summation order
    assert rep["final_rmse"] < 1e-9


def test_oracle_outlier_gating_matches_reference():
    """EdgeSet::setOutlierThreshold + updateEdges at the end of optimize() (src/optimisable_graph.hpp:415,603-640,
    src/cuda_graph_optimisation.cpp:150-151): optimize(5), optimize(5) again on the unmodified reference
    (tests/golden/ref_small_outlier.*) against the oracle -- chi2 of both runs, outlier masks and counts."""
    rep = golden_util.check_outlier_case("small_outlier", make_oracle)
    print(rep)
    assert rep["outliers_a"] > 0 and rep["final_rmse"] < 1e-9
