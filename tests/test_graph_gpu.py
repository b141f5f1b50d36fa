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


def test_cov_to_information(engine):
    """information = covariance.inverse() of matchNodePairICP (MapNode.cpp:638-651) and the constant
    override (MapGraph.cpp:1427-1429).  Inputs: the covariances zhicp really returns (the frozen oracle outputs on the
    reference's node clouds) plus random SPD matrices over 12 orders of magnitude and singular ones."""
    import os
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "zhicp_nodes.npz"))
    c = d["oracle"]["cov"].reshape(-1, 3, 3)
    real = np.stack([c[:, 0, 0], c[:, 0, 1], c[:, 0, 2], c[:, 1, 1], c[:, 1, 2], c[:, 2, 2]], 1)
    rng = np.random.default_rng(7)
    A = rng.normal(0, 1, (4000, 3, 3)) * (10.0 ** rng.uniform(-6, 6, (4000, 1, 1)))
    spd = A @ A.transpose(0, 2, 1)
    rnd = np.stack([spd[:, 0, 0], spd[:, 0, 1], spd[:, 0, 2], spd[:, 1, 1], spd[:, 1, 2], spd[:, 2, 2]], 1)
    sing = np.array([[0, 0, 0, 0, 0, 0], [1, 0, 0, 0, 0, 0.0]])   # singular whatever the scale
    cov6 = np.concatenate([real, rnd, sing])
    for scale in (1.0, 2.5):
        got, ok = engine.cov_to_information(cov6, scale=scale)
        want = po.cov_to_information(cov6, scale=scale)
        n_ok = real.shape[0] + rnd.shape[0]
        assert np.array_equal(got[:n_ok], want[:n_ok])          # same formula, same order: the same bits
        assert ok[:n_ok].all() and not ok[n_ok:].any()
    # it IS the inverse: info * cov = I (well-conditioned part), and it is symmetric like Eigen's result
    got, _ = engine.cov_to_information(real)
    full = np.stack([real[:, [0, 1, 2]], real[:, [1, 3, 4]], real[:, [2, 4, 5]]], 1)
    assert np.allclose(got @ full, np.eye(3), atol=1e-6)
    assert np.allclose(got, np.linalg.inv(full), rtol=1e-6, atol=0)
    # the constant-covariance mode: the graph's m_cov_loopnode = diag(10, 10, 10) is handed over AS the information
    got, ok = engine.cov_to_information(cov6, const_diag=[10.0, 10.0, 10.0])
    assert np.array_equal(got, np.broadcast_to(np.diag([10.0, 10.0, 10.0]), got.shape))
    assert np.array_equal(got, po.cov_to_information(cov6, const_diag=[10.0, 10.0, 10.0]))
