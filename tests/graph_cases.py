"""Seeded inputs for the loop-closure gating / consensus tests (shared by the CPU and GPU test files)."""
import numpy as np


def node_table(N=400, seed=11):
    """A trajectory of submap nodes wandering through a 40 m x 40 m area with loops, node-frame bounding boxes like
    CMapNode keeps (m_ld_corner / m_ru_corner, MapNode.cpp:152-155), and growing pose covariances."""
    rng = np.random.default_rng(seed)
    th = np.cumsum(rng.normal(0, 0.25, N))
    step = rng.uniform(0.3, 1.2, N)
    x = np.cumsum(step * np.cos(th)) % 40.0
    y = np.cumsum(step * np.sin(th)) % 40.0
    pose = np.stack([x, y, th], 1)
    ld = -rng.uniform(0.5, 8.0, (N, 2))
    ru = rng.uniform(0.5, 8.0, (N, 2))
    ld[::37] = 1e6; ru[::37] = -1e6            # a node that never got a point keeps MAX_VALUE / MIN_VALUE corners
    cov = np.abs(rng.normal(0, 1, (N, 2))) * np.linspace(1e-4, 0.05, N)[:, None]
    return pose, ld, ru, cov


def matched_sets(S=300, seed=12):
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, 12, S)
    sizes[:8] = [0, 1, 2, 3, 3, 4, 11, 20]
    off = np.zeros(S + 1, np.int32); off[1:] = np.cumsum(sizes)
    tot = int(off[-1])
    centre = np.repeat(rng.uniform(-20, 20, (S, 3)) * [1, 1, 0.1], sizes, axis=0)
    poses = centre + rng.normal(0, 0.05, (tot, 3))
    out = rng.random(tot) < 0.15                 # mismatches far from the consensus
    poses[out, :2] += rng.normal(0, 2.0, (int(out.sum()), 2))
    good = rng.uniform(0.8, 1.0, tot).astype(np.float32)
    return poses, good, off
