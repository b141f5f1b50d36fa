"""Builds the loop-closure search workload of BASELINE config 3 from a CARMEN log (shared by the GPU integration test):
nodes of K consecutive scans with ground-truth poses, their clouds, the node table, candidate pairs and initial guesses."""
import numpy as np


def rel_pose(a, b):
    """a.ominus(b): pose b expressed in the frame of pose a (src/zhpsm/point.cpp:46-52)."""
    c, s = np.cos(a[2]), np.sin(a[2])
    d = b[:2] - a[:2]
    return np.array([c * d[0] + s * d[1], -s * d[0] + c * d[1], b[2] - a[2]])


def oplus(a, t):
    c, s = np.cos(a[2]), np.sin(a[2])
    return np.array([a[0] + t[0] * c - t[1] * s, a[1] + t[0] * s + t[1] * c, a[2] + t[2]])


def make_nodes(g, K=15, stride=15, max_nodes=70):
    S = g["scans"].shape[0]
    nodes = []
    for s0 in range(0, S - K, stride):
        lp = g["laser_pose"][s0:s0 + K]
        traj = np.stack([rel_pose(lp[0], p) for p in lp])
        nodes.append(dict(start=s0, scans=g["scans"][s0:s0 + K], traj=traj, root=lp[0].copy()))
        if len(nodes) >= max_nodes:
            break
    return nodes


def node_table(nodes, clouds):
    pose = np.stack([nd["root"] for nd in nodes])
    ld = np.stack([[c[0].min(), c[1].min()] if c[0].size else [1e6, 1e6] for c in clouds]).astype(np.float64)
    ru = np.stack([[c[0].max(), c[1].max()] if c[0].size else [-1e6, -1e6] for c in clouds]).astype(np.float64)
    return pose, ld, ru


def candidates(flags, adj_win=2):
    """(cur, ref) pairs the reference would hand to the ICP: both gates pass and ref is outside the adjacency window
    of older nodes (MapGraph.cpp:1342-1352)."""
    N = flags.shape[0]
    out = []
    for i in range(N):
        for j in range(0, max(0, i - adj_win)):
            if flags[i, j] == 3:
                out.append((i, j))
    return np.array(out, np.int32).reshape(-1, 2)


def guesses(nodes, cand, G=8, seed=20261104):
    rng = np.random.default_rng(seed)
    init = np.zeros((len(cand), G, 3))
    for k, (i, j) in enumerate(cand):
        gt = rel_pose(nodes[j]["root"], nodes[i]["root"])
        init[k] = gt + rng.normal(0, 1, (G, 3)) * [0.15, 0.15, np.deg2rad(5)]
        init[k, 0] = gt
    return init
