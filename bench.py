#!/usr/bin/env python
"""bench.py -- PL-ICP scan-pair matches/s on synthetic 1080-beam scans (BASELINE.json configs[3]).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference ...                     (the CPU restatement on the host cores, rank 0 only)

One step = one pass of the hot path over one batch: D2D restore of the raw scan pool -> in-place median filter
(CCanonicalMatcher::median_filter) -> batched sm_icp (zhsm_plicp_batch) -> [N>1] NCCL all-gather of the 128-byte
result records.  Weak scaling: every rank owns `--pairs` independent pairs (Philox counter offset by rank), no
data-path collective besides the result all-gather the north star names.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "plicp_scan_pair_matches_per_sec"
UNIT = "pairs/s"
N_BEAMS = 1080
FOV_DEG = 270.0
MAX_RANGE_CM = 3000
SEED = 20261104
# dram__bytes_read.sum + dram__bytes_write.sum of plicp2_kernel per pair (profiles/r01c_*: 175.5 MB + 30.1 MB / 20000 pairs)
NCU_DRAM_BYTES_PER_PAIR = (348.270080e6 + 39.143680e6) / 20000  # profiles/r01j_plicp2_T192_raw_metrics.txt


def make_laser(zh):
    lz = zh.Laser()
    lz.n = N_BEAMS
    lz.fi_min = np.float32(-FOV_DEG / 2 * np.pi / 180)
    lz.dfi = np.float32(FOV_DEG * np.pi / 180 / (N_BEAMS - 1))
    lz.min_range_cm = 10
    lz.max_range_cm = MAX_RANGE_CM
    lz.min_valid_points = N_BEAMS // 5   # the reference tables' ratio (PolarParameter.h:70-83)
    lz.scan_window = N_BEAMS // 10
    return lz


def workload_config(pairs, n_gpus):
    return {"workload": "synthetic 1080-beam Hokuyo-style scans (SURVEY 8d C4: room+boxes ray cast, 1 cm noise, "
                        "1 mm quantisation, 1% dropouts), PL-ICP first guess 0",
            "beams": N_BEAMS, "fov_deg": FOV_DEG, "max_range_m": MAX_RANGE_CM / 100,
            "pairs_per_gpu": pairs, "pairs_total": pairs * n_gpus, "seed": SEED,
            "l2": "inputs (2*pairs*1080*4 B per step) are far larger than the 126 MB L2",
            "parallelism": f"shard{n_gpus}"}


def algorithmic_flops(sample_scans_cm, lz, iterations_sum, npairs):
    """F_pair = sum_iter [ n_s*(5*Wbar + 126) + 500 ] + 30*n with Wbar = mean possible_interval width."""
    n = lz.n
    r = sample_scans_cm[1::2].astype(np.float64) / 100.0           # sens scans
    valid = (r > lz.min_range_cm / 100.0) & (r < lz.max_range_cm / 100.0)
    n_s = valid.sum(1).mean()
    ang_res = (float(lz.dfi) * (n - 1)) / n
    rho = np.where(valid, r, np.nan)
    R = np.ceil((np.deg2rad(45.0) + np.arctan(0.5 / rho)) / ang_res)
    c = np.arange(n)[None, :]
    W = np.minimum(n - 1, c + R) - np.maximum(0, c - R) + 1
    Wbar = float(np.nanmean(W))
    f_iter = n_s * (5.0 * Wbar + 126.0) + 500.0
    total = iterations_sum * f_iter + npairs * 30.0 * n
    return total, {"n_s": float(n_s), "Wbar": Wbar, "flop_per_iteration": float(f_iter)}


class ClockSampler:
    """nvidia-smi clocks line of B200_PROFILING.md, sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={gpu_index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f:
            t = [x.strip() for x in line.split(",")]
            if len(t) < 9:
                continue
            try:
                sm.append(float(t[1])); mx.append(float(t[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), t[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def gen_host_sample(zh, lz, npairs, first_pair=0, device=0):
    """Synthetic pairs generated by the device-side generator (input data only), returned as a host array."""
    import torch
    eng = zh.Engine(device)
    eng.set_laser(lz)
    dev = torch.device(f"cuda:{device}")
    scans = torch.empty((2 * npairs, lz.n), dtype=torch.float32, device=dev)
    truth = torch.empty((npairs, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(scans, truth, first_pair, npairs, seed=SEED)
    eng.sync()
    out = scans.cpu().numpy()
    eng.close()
    return out


def cpu_time_sample(raw_scans, lz, nthreads):
    """The CPU restatement (oracle 'port') on a bounded sample: median filter + sm_icp per pair, all host threads."""
    from oracle import pyoracle as po
    P = raw_scans.shape[0] // 2
    work = raw_scans.copy()
    ri = np.arange(0, 2 * P, 2, dtype=np.int32)
    si = ri + 1
    t0 = time.perf_counter()
    po.median_filter_batch(work, nthreads)
    res = po.zhcsm_fmatch_batch(work, ri, si, lz.fi_min, lz.dfi, lz.min_range_cm, lz.max_range_cm, nthreads=nthreads)
    dt = time.perf_counter() - t0
    return P / dt, dt, res


def run_reference(args):
    """--impl reference: CSM is not vendored, so the reference's CPU path is the oracle port (kind 'port')."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import laser_slam_b200 as zh
    lz = make_laser(zh)
    cores = os.cpu_count() or 1
    P = args.ref_pairs
    raw = gen_host_sample(zh, lz, P)
    for _ in range(args.warmup):
        cpu_time_sample(raw[: 2 * max(256, P // 8)], lz, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_time_sample(raw, lz, cores)
    dt = time.perf_counter() - t0
    value = P * args.steps / dt
    sample = f"{P} pairs per step of the same synthetic workload, median filter + sm_icp restatement, {cores} pthreads"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.pairs, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def other_paths(zh, eng, dev):
    """Kernel-only rates (CUDA events inside the library) of the other two matchers on the committed fixtures:
    zhpsm on fr079 (BASELINE config 2) replicated 8x, zhicp on the reference's node clouds x 64 guesses (config 3)."""
    import torch
    from laser_slam_b200 import engine as E
    sys.path.insert(0, ROOT)
    from tests.support import load_log, make_laser as fixture_laser
    out = {}
    g = load_log("fr079")
    eng.set_laser(fixture_laser(g))
    scans = torch.from_numpy(g["scans"].copy()).to(dev)
    S, n = scans.shape
    _, bad, seg = eng.preprocess(scans)
    ri = torch.arange(0, S - 1, dtype=torch.int32, device=dev).repeat(8)
    ai = ri + 1
    res = torch.zeros((ri.shape[0], 128), dtype=torch.uint8, device=dev)
    for name, algo in (("pm_psm", E.ALGO_PSM), ("pm_icp", E.ALGO_PM_ICP)):
        for _ in range(2):
            eng.psm(scans, bad, seg, ri, ai, algo=algo, results=res)
            eng.sync()
        out[name] = {"jobs_per_s": ri.shape[0] / (eng.last_kernel_ms() * 1e-3), "jobs": int(ri.shape[0]), "beams": int(n),
                     "workload": "fr079 consecutive pairs x 8"}
    d = np.load(os.path.join(ROOT, "tests", "golden", "zhicp_nodes.npz"))
    G, rep = 64, 4
    gx, gy, gt = np.meshgrid(np.linspace(-0.3, 0.3, 4), np.linspace(-0.3, 0.3, 4), np.linspace(-0.15, 0.15, 4), indexing="ij")
    grid = np.stack([gx.ravel(), gy.ravel(), gt.ravel()], 1)
    init = np.tile((d["init"][:, None, :] + grid[None]).reshape(-1, 3), (rep, 1))
    cri = np.tile(np.repeat(d["pairs"][:, 0], G), rep).astype(np.int32)
    cci = np.tile(np.repeat(d["pairs"][:, 1], G), rep).astype(np.int32)
    pts = torch.from_numpy(d["pts"]).to(dev)
    dri, dci, dinit = torch.from_numpy(cri).to(dev), torch.from_numpy(cci).to(dev), torch.from_numpy(init).to(dev)
    res = torch.zeros((cri.size, 128), dtype=torch.uint8, device=dev)
    for _ in range(2):
        eng.icp(pts, d["off"], d["cnt"], dri, dci, dinit, results=res)
        eng.sync()
    out["zhicp"] = {"jobs_per_s": cri.size / (eng.last_kernel_ms() * 1e-3), "jobs": int(cri.size),
                    "workload": "63 node-cloud pairs of src/zhicp/data{,2} x 64 initial guesses x 4"}
    return out


def init_ranks():
    """world, rank, local rank, device; brings NCCL up (N>1) with stdout kept for the one JSON line."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device(f"cuda:{local}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version banner (and anything NCCL_DEBUG asks for) to stdout while the communicator comes
        # up: stdout is for the one JSON line, so file descriptor 1 points at stderr until the first collective is done
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    return world, rank, local, dev


def run_ours(args):
    import torch
    import torch.distributed as dist
    import laser_slam_b200 as zh
    from laser_slam_b200 import shard

    world, rank, local, dev = init_ranks()
    eng = zh.Engine(local)
    lz = make_laser(zh)
    eng.set_laser(lz)
    P, n = args.pairs, lz.n
    # a dedicated non-default stream: its handle is non-null, so the engine launches on the very stream the
    # torch CUDA events below are recorded on (a null stream pointer would select the handle's own stream)
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0

    raw = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
    work = torch.empty_like(raw)
    truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(raw, truth, rank * P, P, seed=SEED, stream=stream)
    ri = torch.arange(0, 2 * P, 2, dtype=torch.int32, device=dev)
    si = ri + 1
    res = torch.zeros((P, 128), dtype=torch.uint8, device=dev)
    gathered = [None]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    kern_events = []

    def step(record=False):
        work.copy_(raw)
        eng.median_filter(work, stream=stream)
        e0 = e1 = None
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        eng.plicp(work, ri, si, results=res, stream=stream)
        if record:
            e1.record()
            kern_events.append((e0, e1))
        if world > 1:
            gathered[0] = shard.all_gather_records(res, world * P)

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(local) if rank == 0 else None
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record()
    for _ in range(args.steps):
        step(record=True)
    t_end.record()
    barrier()
    ms = t_start.elapsed_time(t_end)
    launches = eng.launch_count() - launches0 + args.steps  # + the D2D restore copy kernel per step
    clk = clocks.stop() if clocks else None
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kern_events]))
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    value = world * P * args.steps / (ms * 1e-3)

    # ---------------- end to end through the C ABI with HOST buffers (pinned), copies inside the timed region
    chunk = min(P, args.e2e_chunk)
    host_raw = torch.empty((2 * P, n), dtype=torch.float32, pin_memory=True)
    host_raw.copy_(raw)
    host_np = host_raw.numpy()
    host_res = torch.empty((P, 128), dtype=torch.uint8, pin_memory=True)
    res_np = host_res.numpy().view(zh.RESULT_DTYPE).reshape(-1)
    ri_h = np.arange(0, 2 * chunk, 2, dtype=np.int32)
    si_h = ri_h + 1

    # Two handles, one per host thread ("one handle = one GPU + one stream + its staging buffers: use one
    # handle per host thread"): while one thread's chunk is being matched, the other thread's chunk is on the PCIe bus.
    # ctypes releases the GIL for the duration of a call, so plain Python threads are enough.
    import concurrent.futures as cf
    eng2 = zh.Engine(local)
    eng2.set_laser(lz)
    engines = [eng, eng2]
    pool = cf.ThreadPoolExecutor(max_workers=2)
    chunks = [(c0, min(P, c0 + chunk)) for c0 in range(0, P, chunk)]

    def e2e_worker(w):
        torch.cuda.set_device(local)
        for c0, c1 in chunks[w::2]:
            k = c1 - c0
            engines[w].plicp(host_np[2 * c0: 2 * c1], ri_h[:k], si_h[:k], results=res_np[c0:c1], filter_scans=True)

    def e2e_step():
        for f in [pool.submit(e2e_worker, w) for w in range(2)]:
            f.result()

    e2e_step()  # warm the staging buffers
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    barrier()
    e2e_dt = time.perf_counter() - t0
    te = torch.tensor([e2e_dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = world * P * args.e2e_steps / float(te.item())
    h2d = 2 * P * n * 4 + 2 * P * 4
    d2h = P * 128

    if rank == 0:
        r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        # parity of the device path vs the host-buffer path on the same inputs (bit-identical records)
        same = bool((r["corr_hash"] == res_np["corr_hash"]).all() and (r["iterations"] == res_np["iterations"]).all())
        it_sum = int(r["iterations"].sum())
        ns = min(P, 2048)
        flops, fl_detail = algorithmic_flops(host_np[: 2 * ns], lz, it_sum, P)
        peak64 = eng.fp_peak_tflops(64)
        peak32 = eng.fp_peak_tflops(32)
        achieved = flops / (kern_ms * 1e-3) / 1e12
        hbm_bytes = P * (2 * n * 4 + 128)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        cpu = None
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            cp = min(P, args.cpu_pairs)
            v, dt, ores = cpu_time_sample(host_np[: 2 * cp], lz, cores)
            g = r[:cp]
            both = (g["valid"] == 1) & (ores["valid"] == 1)
            d = np.abs(g["x"] - ores["x"])
            ok = (d[:, 0] <= 1e-4) & (d[:, 1] <= 1e-4) & (d[:, 2] <= 1e-5)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"first {cp} pairs of the same batch, median filter + sm_icp restatement, {dt:.1f} s",
                   "parity_on_sample": {"valid_eq": float((g["valid"] == ores["valid"]).mean()),
                                        "pose_within_1e-4m_1e-5rad": float(ok[both].mean()) if both.any() else None,
                                        "iterations_eq": float((g["iterations"] == ores["iterations"]).mean()),
                                        "corr_hash_eq": float((g["corr_hash"] == ores["corr_hash"])[both].mean()) if both.any() else None}}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(P, world),
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": args.e2e_steps, "api": "zhsm_plicp_batch(host buffers, ZHSM_FILTER_SCANS), 2 handles on 2 host threads, chunks of %d pairs" % chunk,
                        "device_vs_host_path_identical": same},
                "gpu_launches": int(launches),
                "roofline": {"bound": "fp64_cuda_core", "achieved": achieved, "peak": peak64, "unit": "TFLOP/s",
                             "frac": achieved / peak64 if peak64 > 0 else None,
                             "traffic": NCU_DRAM_BYTES_PER_PAIR * P,
                             "traffic_source": "ncu --set full dram__bytes_read+write of a 20k-pair launch "
                                               "(profiles/), scaled by pairs per launch",
                             "kernel": "plicp2_kernel<%d>" % (128 if N_BEAMS <= 640 else 192 if N_BEAMS <= 1472 else 256 if N_BEAMS <= 2304 else 512), "kernel_ms": kern_ms,
                             "peak_source": "DFMA micro-benchmark measured in this run (zhsm_fp_peak_tflops)",
                             "fp32_peak": peak32, "frac_of_fp32_peak": achieved / peak32 if peak32 > 0 else None,
                             "algorithmic_flops_per_launch": flops, **fl_detail,
                             "iterations_per_pair": it_sum / P,
                             "hbm": {"achieved_gbs": hbm_bytes / (kern_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                                     "frac": hbm_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak,
                                     "algorithmic_bytes_per_launch": hbm_bytes}},
                "results": {"valid_frac": float((r["valid"] == 1).mean()),
                            "median_abs_err_vs_truth": np.median(np.abs(r["x"] - truth.cpu().numpy()), 0).tolist()}}
        if cpu:
            line["cpu_baseline"] = cpu
        if world == 1:
            try:
                line["other_paths"] = other_paths(zh, eng, dev)
            except Exception as e:  # the headline line must not depend on the side measurements
                line["other_paths"] = {"error": repr(e)}
            eng.set_laser(lz)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


C5_BEAMS, C5_STARTS, C5_SEED = 2048, 16, 20261105


def run_c5(args):
    """--workload c5 (BASELINE config 5): multi-start PL-ICP for loop closure.  Per GPU `--pairs` scan
    pairs (default 7812 = 62 500 / 8) of 2048-beam 360-degree scans with a true offset up to 0.5 m / 20 degrees, 16 first
    guesses per pair = noisy odometry (+) {-0.3, +0.3}^2 m x {-10, -3.3, +3.3, +10} degrees.  A step = median filter ->
    one zhsm_plicp_batch over all (pair, start) jobs -> zhsm_best_of_starts (lowest mean error) on the device ->
    [N>1] NCCL all-gather of the best record per pair (pose + covariance, 128 bytes).  Pairs are sharded by pair, all
    starts of a pair on one GPU."""
    import torch
    import torch.distributed as dist
    import laser_slam_b200 as zh
    from laser_slam_b200 import engine as E
    from laser_slam_b200 import shard
    world, rank, local, dev = init_ranks()
    n, G = C5_BEAMS, C5_STARTS
    P = args.pairs if args.pairs != 1_000_000 else 7812
    PM_PI = 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = MAX_RANGE_CM; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    eng = zh.Engine(local)
    eng.set_laser(lz)
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    raw = torch.empty((2 * P, n), dtype=torch.float32, device=dev)
    work = torch.empty_like(raw)
    truth = torch.empty((P, 3), dtype=torch.float64, device=dev)
    eng.synth_pairs(raw, truth, rank * P, P, dmax_xy_m=0.5, dmax_th_rad=np.deg2rad(20.0), seed=C5_SEED, stream=stream)
    torch.cuda.synchronize()
    tr = truth.cpu().numpy()
    rng = np.random.default_rng(C5_SEED + rank)
    odo = tr + rng.normal(0, 1, (P, 3)) * [0.1, 0.1, np.deg2rad(3.0)]
    grid = np.array([[dx, dy, np.deg2rad(dt)] for dx in (-0.3, 0.3) for dy in (-0.3, 0.3) for dt in (-10, -3.3, 3.3, 10)])
    guess_h = (odo[:, None, :] + grid[None]).reshape(-1, 3)
    ri_h = np.repeat(np.arange(P, dtype=np.int32) * 2, G)
    ri, si, guess = (torch.from_numpy(a).to(dev) for a in (ri_h, ri_h + 1, guess_h))
    res = torch.zeros((P * G, 128), dtype=torch.uint8, device=dev)
    out = [None, None]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    kern_events = []

    def step(record=False):
        work.copy_(raw)
        eng.median_filter(work, stream=stream)
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        eng.plicp(work, ri, si, first_guess=guess, results=res, stream=stream)
        if record:
            e1.record()
            kern_events.append((e0, e1))
        best, rec = eng.best_of_starts(res, P, G, criterion=E.BEST_MEAN_ERROR, stream=stream)
        out[0], out[1] = best, (shard.all_gather_records(rec, world * P) if world > 1 else rec)

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(local) if rank == 0 else None
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        step(record=True)
    t1.record()
    barrier()
    ms = t0.elapsed_time(t1)
    launches = eng.launch_count() - launches0 + args.steps
    clk = clocks.stop() if clocks else None
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kern_events]))
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    value = world * P * G * args.steps / (ms * 1e-3)

    # end to end with HOST buffers: scans, job lists and first guesses go up, the best record per pair comes back
    host_raw = torch.empty((2 * P, n), dtype=torch.float32, pin_memory=True)
    host_raw.copy_(raw)
    host_np = host_raw.numpy()
    si_h = ri_h + 1

    def e2e_step():
        r_dev = torch.empty((P * G, 128), dtype=torch.uint8, device=dev)
        eng.plicp(host_np, ri_h, si_h, first_guess=guess_h, results=r_dev, filter_scans=True, stream=stream)
        _, rec = eng.best_of_starts(r_dev, P, G, criterion=E.BEST_MEAN_ERROR, stream=stream)
        if world > 1:
            rec = shard.all_gather_records(rec, world * P)
        return rec.cpu()

    e2e_step()
    barrier()
    w0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        rec_h = e2e_step()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - w0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * P * G * args.e2e_steps / float(e2e_s.item())

    r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(P, G)
    b = out[1].cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)[rank * P:(rank + 1) * P]
    err = np.abs(b["x"] - tr); err[:, 2] = np.abs((err[:, 2] + np.pi) % (2 * np.pi) - np.pi)
    ok = (b["valid"] == 1) & (err[:, 0] < 0.02) & (err[:, 1] < 0.02) & (err[:, 2] < np.deg2rad(0.5))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    sample = work[: 2 * min(P, 256)].cpu().numpy()
    flops, fdet = algorithmic_flops(sample, lz, float(r["iterations"].sum()), P * G)
    fp64_peak = eng.fp_peak_tflops(64)
    ach = flops / (kern_ms * 1e-3) / 1e12
    line = {"metric": "plicp_multistart_matches_per_sec", "value": value, "unit": "matches/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "synthetic 2048-beam wide-FOV scans, multi-start PL-ICP for loop closure (SURVEY 8d C5): 16 "
                                   "first guesses per pair around a noisy odometry, best-of-16 on the device, best record "
                                   "per pair all-gathered", "beams": n, "starts": G, "pairs_per_gpu": P,
                       "pairs_total": P * world, "seed": C5_SEED, "parallelism": f"shard{world} by pair",
                       "l2": "inputs (2*pairs*2048*4 B per step = %.0f MB) are larger than the 126 MB L2" % (2 * P * n * 4 / 1e6)},
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": "matches/s", "steps": args.e2e_steps,
                    "h2d_bytes_per_step": int(2 * P * n * 4 + P * G * (8 + 24)), "d2h_bytes_per_step": int(P * world * 128),
                    "api": "zhsm_plicp_batch(host scans + jobs + first guesses, ZHSM_FILTER_SCANS) -> zhsm_best_of_starts -> best records to the host"},
            "gpu_launches": launches,
            "roofline": {"bound": "fp64_cuda_core", "achieved": ach, "peak": fp64_peak, "unit": "TFLOP/s", "frac": ach / fp64_peak,
                         "traffic": None, "kernel": "plicp2_kernel<256>", "kernel_ms": kern_ms,
                         "peak_source": "DFMA micro-benchmark measured in this run (zhsm_fp_peak_tflops)",
                         "algorithmic_flops_per_launch": flops, **fdet,
                         "iterations_per_start": float(r["iterations"].mean())},
            "results": {"valid_frac": float((r["valid"] == 1).mean()),
                        "pairs_with_best_start_within_2cm_0.5deg_of_truth": float(ok.mean())}}
    if world == 1 and not args.no_cpu:
        from oracle import pyoracle as po
        K = 512
        cores = os.cpu_count() or 1
        fm = work[: 2 * (K // G)].cpu().numpy()
        w0 = time.perf_counter()
        o = po.zhcsm_fmatch_batch(fm, ri_h[:K], si_h[:K], lz.fi_min, lz.dfi, lz.min_range_cm, lz.max_range_cm,
                                  first_guess=guess_h[:K], nthreads=cores)
        dt = time.perf_counter() - w0
        rr = r.reshape(-1)[:K]
        both = (rr["valid"] == 1) & (o["valid"] == 1)
        d = np.abs(rr["x"] - o["x"])
        line["cpu_baseline"] = {"value": K / dt, "unit": "matches/s", "cores": cores, "kind": "port",
                                "sample": "the first %d (pair, start) jobs of the same batch, sm_icp restatement, %.2f s" % (K, dt),
                                "parity_on_sample": {"valid_eq": float((rr["valid"] == o["valid"]).mean()),
                                                     "pose_within_1e-4m_1e-5rad": float(((d[:, 0] <= 1e-4) & (d[:, 1] <= 1e-4) & (d[:, 2] <= 1e-5))[both].mean()),
                                                     "iterations_eq": float((rr["iterations"] == o["iterations"])[both].mean()),
                                                     "corr_hash_eq": float((rr["corr_hash"] == o["corr_hash"])[both].mean())}}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=1_000_000, help="pairs per GPU per step (configs[3]: 1M)")
    ap.add_argument("--e2e-steps", type=int, default=1)
    ap.add_argument("--e2e-chunk", type=int, default=125_000)
    ap.add_argument("--cpu-pairs", type=int, default=100_000, help="cpu_baseline sample (pairs)")
    ap.add_argument("--ref-pairs", type=int, default=20_000, help="--impl reference: pairs per step")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--workload", choices=["c4", "c5"], default="c4",
                    help="c4 (default): the headline metric, 1080-beam odometry pairs; c5: 2048-beam multi-start loop closure")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "c5":
        run_c5(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
