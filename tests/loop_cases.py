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


# ---- BASELINE config 3 exactly as SURVEY 8d states it -----------------------------------------------------------------
C3_SEED, C3_G = 20261104, 64


def pairs_c3(nodes, flags, adj_win=2):
    """Config 3's pair list: the consecutive node pairs (k, k+1) plus the loop candidates (GT poses within 2 m =
    isLoopyArea, bbox overlap >= 0.4 = isOverlappedArea, outside the adjacency window).  Rows are (cur, ref)."""
    N = len(nodes)
    consec = np.stack([np.arange(1, N), np.arange(0, N - 1)], 1).astype(np.int32)
    return np.concatenate([consec, candidates(flags, adj_win)]).astype(np.int32)


def guesses_c3(nodes, pairs, G=C3_G, seed=C3_SEED):
    """64 initial guesses per pair: guess 0 = the GT relative pose, guesses 1..63 = guess 0 (+) U(+-0.5 m, +-0.5 m,
    +-15 deg), Philox key = seed, counter = pair * 64 + g."""
    init = np.zeros((len(pairs), G, 3))
    half = np.array([0.5, 0.5, np.deg2rad(15.0)])
    for k, (i, j) in enumerate(pairs):
        gt = rel_pose(nodes[j]["root"], nodes[i]["root"])
        init[k, 0] = gt
        for gidx in range(1, G):
            u = np.random.Generator(np.random.Philox(key=seed, counter=[k * G + gidx, 0, 0, 0])).uniform(-1.0, 1.0, 3)
            init[k, gidx] = oplus(gt, u * half)
    return init


def build_c3(g, clouds_of, gate, K=15, stride=5, max_nodes=400):
    """nodes -> clouds (clouds_of(node) -> (x, y)) -> node table -> gate(pose, ld, ru, cur) -> pairs + guesses."""
    nodes = make_nodes(g, K=K, stride=stride, max_nodes=max_nodes)
    clouds = [clouds_of(nd) for nd in nodes]
    pose, ld, ru = node_table(nodes, clouds)
    flags = gate(pose, ld, ru, np.arange(len(nodes)))
    pairs = pairs_c3(nodes, flags)
    return nodes, clouds, pose, flags, pairs, guesses_c3(nodes, pairs)
