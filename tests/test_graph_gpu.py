"""Loop-closure gating and consensus kernels (zhsm_gate_candidates / zhsm_available_area / zhsm_consensus, SURVEY
8f-1) against the CPU restatement oracle/graph_oracle.c on the same seeded inputs.  Flags, keep masks and chosen
indices are index work: they must be identical; the weighted pose is float arithmetic in the reference's order:
identical bits expected, 1e-6 allowed."""
import numpy as np
import pytest

from oracle import pyoracle as po
from tests.graph_cases import matched_sets, node_table

pytestmark = pytest.mark.gpu


def test_gate_candidates(engine):
    pose, ld, ru, cov = node_table(N=700, seed=3)
    N = pose.shape[0]
    cur = np.arange(N, dtype=np.int32)
    for cov_xy, beta in ((None, 225.0), (cov, 225.0), (cov, 25.0)):
        want = po.gate_candidates(pose, ld, ru, cur, cov_xy=cov_xy, beta=beta)
        got = engine.gate_candidates(pose, ld, ru, cur, cov_xy=cov_xy, beta=beta)
        assert np.array_equal(got, want), np.argwhere(got != want)[:5]
    # a subset of current nodes (the incremental use: the newest nodes against the whole graph)
    sub = np.array([N - 1, N - 2, 5], np.int32)
    assert np.array_equal(engine.gate_candidates(pose, ld, ru, sub), po.gate_candidates(pose, ld, ru, sub))


def test_available_area(engine):
    rng = np.random.default_rng(4)
    M = 5000
    trans = rng.normal(0, 1, (M, 3)) * [1.5, 1.5, 0.7]
    cov = np.abs(rng.normal(0, 0.05, (M, 2)))
    for c in (None, cov):
        assert np.array_equal(engine.available_area(trans, c), po.available_area(trans, c))
    assert 0.1 < po.available_area(trans).mean() < 0.9


def test_consensus(engine):
    poses, good, off = matched_sets(S=2000, seed=21)
    k0, p0, i0 = po.consensus(poses, good, off)
    k1, p1, i1 = engine.consensus(poses, good, off)
    assert np.array_equal(k0, k1) and np.array_equal(i0, i1)
    assert np.allclose(p0, p1, rtol=0, atol=1e-6)
    assert (p0 == p1).mean() > 0.999
