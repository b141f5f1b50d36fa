#!/usr/bin/env python
"""bench.py -- PL-ICP scan-pair matches/s on synthetic 1080-beam scans (BASELINE.json configs[3], "C4"), with the other
four BASELINE configs measured beside it in the same JSON line (`configs`: C1, C2, C3, C5) when N = 1.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference ...                     (the CPU restatement on the host cores, rank 0 only)
  python bench.py --workload c5 ...                        (config 5 as a line of its own, also under torchrun)

One C4 step = one pass of the hot path over one batch: D2D restore of the raw scan pool -> in-place median filter
(CCanonicalMatcher::median_filter) -> batched sm_icp (zhsm_plicp_batch) -> [N>1] NCCL all-gather of the 128-byte
result records, issued on a side stream so that it overlaps the next step's kernel.  Weak scaling: every
rank owns `--pairs` independent pairs (Philox counter offset by rank), no data-path collective besides that gather.
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
# dram__bytes_read.sum + dram__bytes_write.sum of one plicp2_kernel<192> launch over 20 000 pairs, from the committed
# capture named here (tools/ncu_summary.py prints both numbers)
NCU_DRAM_CAPTURE = "profiles/r03f_plicp2_T192_raw_metrics.txt"
NCU_DRAM_BYTES_PER_PAIR = (352.156416e6 + 43.235840e6) / 20000


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


def plicp_flops(sens_scans_cm, n, dfi, min_range_cm, max_range_cm, iterations_sum, njobs):
    """F_pair = sum_iter [ n_s*(5*Wbar + 126) + 500 ] + 30*n with Wbar = mean possible_interval width (the
    exhaustive-in-window formulation) and the MEASURED sum of sm_result.iterations."""
    r = sens_scans_cm.astype(np.float64) / 100.0
    valid = (r > min_range_cm / 100.0) & (r < max_range_cm / 100.0)
    n_s = valid.sum(1).mean()
    ang_res = (float(dfi) * (n - 1)) / n
    rho = np.where(valid, r, np.nan)
    R = np.ceil((np.deg2rad(45.0) + np.arctan(0.5 / rho)) / ang_res)
    c = np.arange(n)[None, :]
    W = np.minimum(n - 1, c + R) - np.maximum(0, c - R) + 1
    Wbar = float(np.nanmean(W))
    f_iter = n_s * (5.0 * Wbar + 126.0) + 500.0
    total = iterations_sum * f_iter + njobs * 30.0 * n
    return total, {"n_s": float(n_s), "Wbar": Wbar, "flop_per_iteration": float(f_iter)}


def algorithmic_flops(sample_scans_cm, lz, iterations_sum, npairs):
    """(kept for tools/ and tests: pairs interleaved ref, sens)"""
    return plicp_flops(sample_scans_cm[1::2], lz.n, lz.dfi, lz.min_range_cm, lz.max_range_cm, iterations_sum, npairs)


def plicp_executed_flops(records, n_s, n):
    """The FLOPs of the candidates the kernel really evaluates ("a kernel that prunes harder than the window
    reports both"): 5 per executed walk step (zhsm_result.reserved[3], counted by the walks themselves: speculated
    walks, re-searches and the warp-wide windows alike) + the same per-ray and per-iteration terms as the algorithmic
    figure.  (None, None) when the build does not count steps."""
    steps = records["reserved"][:, 3].astype(np.float64)
    if not (steps > 0).any():
        return None, None
    it = float(records["iterations"].sum())
    total = 5.0 * float(steps.sum()) + it * (n_s * 126.0 + 500.0) + records.shape[0] * 30.0 * n
    return total, float(steps.sum() / max(it * n_s, 1.0))


def roofline_block(flops, detail, kern_ms, peak, peak_name, kernel, executed=None, extra=None):
    ach = flops / (kern_ms * 1e-3) / 1e12
    out = {"bound": peak_name, "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak if peak > 0 else None,
           "traffic": None, "kernel": kernel, "kernel_ms": kern_ms,
           "peak_source": "FMA micro-benchmark measured in this run (zhsm_fp_peak_tflops); MEASURED_PEAKS.json holds "
                          "HBM and bf16 tensor peaks only, neither bounds this kernel",
           "algorithmic_flops_per_launch": flops, **(detail or {})}
    if executed is not None and executed[0] is not None:
        ex = executed[0] / (kern_ms * 1e-3) / 1e12
        out["executed"] = {"flops_per_launch": executed[0], "achieved": ex, "frac": ex / peak if peak > 0 else None,
                           "walk_steps_per_ray_and_iteration": executed[1],
                           "note": "5 flops per walk step the kernel executed (speculation and re-searches included) + "
                                   "the per-ray / per-iteration terms of the algorithmic figure"}
    if extra:
        out.update(extra)
    return out


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


def parity_block(g, o):
    """GPU records vs oracle records of the same jobs: north_star's four criteria as fractions."""
    both = (g["valid"] == 1) & (o["valid"] == 1)
    d = np.abs(g["x"] - o["x"])
    ok = (d[:, 0] <= 1e-4) & (d[:, 1] <= 1e-4) & (d[:, 2] <= 1e-5)
    return {"valid_eq": float((g["valid"] == o["valid"]).mean()),
            "pose_within_1e-4m_1e-5rad": float(ok[both].mean()) if both.any() else None,
            "iterations_eq": float((g["iterations"] == o["iterations"]).mean()),
            "corr_hash_eq": float((g["corr_hash"] == o["corr_hash"])[both].mean()) if both.any() else None}


class HostLaser:
    """The C4 geometry without importing the product package (the reference arm maps oracle/ only)."""
    n = N_BEAMS
    fi_min = np.float32(-FOV_DEG / 2 * np.pi / 180)
    dfi = np.float32(FOV_DEG * np.pi / 180 / (N_BEAMS - 1))
    min_range_cm = 10
    max_range_cm = MAX_RANGE_CM


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores.  CSM is not vendored
    under the reference tree, so this is the oracle port (kind "port") with one pthread per core.  Input: the host twin
    of the workload generator (oracle/synth_oracle.c) -- nothing of the product library is loaded by this arm."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle as po
    lz = HostLaser
    cores = os.cpu_count() or 1
    P = args.ref_pairs
    raw, _ = po.synth_pairs(P, lz.n, lz.fi_min, lz.dfi, lz.max_range_cm, seed=SEED, nthreads=cores)
    for _ in range(args.warmup):
        cpu_time_sample(raw[: 2 * max(256, P // 8)], lz, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_time_sample(raw, lz, cores)
    dt = time.perf_counter() - t0
    value = P * args.steps / dt
    sample = (f"{P} pairs per step of the same synthetic workload (host twin of the generator, oracle/synth_oracle.c), "
              f"median filter + sm_icp restatement, {cores} pthreads")
    cfg = workload_config(args.pairs, args.gpus)
    cfg["reference_arm"] = {"pairs_per_step": P, "pairs_run_total": P * args.steps,
                            "note": "a bounded sample of the workload per step; the metric is a rate"}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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


def time_events(fn, reps, warm=1):
    """Mean CUDA-event time (ms) of fn() over reps calls on torch's current stream, after warm untimed calls."""
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


# ======================================================================================================================
# The other BASELINE configs, measured beside the headline when N = 1 (`configs` in the JSON line).  Every block carries
# value (device-resident rate, CUDA events), roofline (SURVEY 8d's algorithmic FLOPs with MEASURED iteration counts over
# the measured FMA peak of the pipe the kernel computes in), e2e (host buffers through the C ABI, copies timed, >= 3
# steps) and cpu_baseline (the compiled reference where it builds -- zhpsm --, the oracle port elsewhere).
# ======================================================================================================================
def _records(zh, t):
    return t.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)


def measure_c1(zh, eng, dev, stream, peaks, args):
    """BASELINE configs[0]: zhcsm PL-ICP consecutive-scan odometry over data/intel-lab.log (CARMEN, 180 beams read +
    1 padded), first guess 0 -- the config where the restart loop of sm_icp fires on most pairs."""
    import torch
    from oracle import pyoracle as po
    from tests.support import load_log, make_laser as fixture_laser
    g = load_log("intel-lab")
    lz = fixture_laser(g)
    eng.set_laser(lz)
    raw_h = np.ascontiguousarray(g["scans"])
    S, n = raw_h.shape
    B = S - 1
    rep = 16   # the log x 16 in one batch: enough CTAs (42 736) for a steady-state rate; the single log is 4.5 waves
    raw = torch.from_numpy(raw_h).to(dev).repeat(rep, 1)
    work = torch.empty_like(raw)
    ri = torch.cat([torch.arange(0, B, dtype=torch.int32, device=dev) + k * S for k in range(rep)])
    si = ri + 1
    res = torch.zeros((B * rep, 128), dtype=torch.uint8, device=dev)

    def step():
        work.copy_(raw)
        eng.median_filter(work, stream=stream)
        eng.plicp(work, ri, si, results=res, stream=stream)
    ms = time_events(step, reps=10, warm=3)
    kern_ms = eng.last_kernel_ms()
    r = _records(zh, res)
    one = r[:B]
    flops, det = plicp_flops(work[1:S].cpu().numpy(), n, lz.dfi, lz.min_range_cm, lz.max_range_cm, float(r["iterations"].sum()), B * rep)
    ex = plicp_executed_flops(r, det["n_s"], n)
    # the log once, as the reference runs it
    ri1, si1 = ri[:B].contiguous(), si[:B].contiguous()
    res1 = torch.zeros((B, 128), dtype=torch.uint8, device=dev)

    def step1():
        work[:S].copy_(raw[:S])
        eng.median_filter(work[:S], stream=stream)
        eng.plicp(work[:S], ri1, si1, results=res1, stream=stream)
    ms1 = time_events(step1, reps=20, warm=3)
    # e2e: host scans + host job lists in, host records out, the median filter inside the call
    ri_h, si_h = np.arange(0, B, dtype=np.int32), np.arange(1, S, dtype=np.int32)
    out_h = np.zeros(B, zh.RESULT_DTYPE)
    eng.plicp(raw_h, ri_h, si_h, results=out_h, filter_scans=True, stream=stream)
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        eng.plicp(raw_h, ri_h, si_h, results=out_h, filter_scans=True, stream=stream)
    e2e_dt = (time.perf_counter() - t0) / args.e2e_steps
    same = bool(np.array_equal(out_h["corr_hash"], one["corr_hash"]) and np.array_equal(out_h["iterations"], one["iterations"]))
    blk = {"workload": "zhcsm PL-ICP consecutive-scan odometry over data/intel-lab.log (2671 pairs, 181 beams), first guess 0",
           "metric": METRIC, "unit": UNIT, "value": B * rep / (ms * 1e-3),
           "value_note": f"the log's 2671 pairs x {rep} in one batch (steady state); the log once: see single_log",
           "single_log": {"pairs": B, "ms": ms1, "value": B / (ms1 * 1e-3)},
           "dtype": "f64", "beams": int(n), "jobs_per_step": int(B * rep), "ms_per_step": ms,
           "iterations_per_pair": float(one["iterations"].mean()), "icp_loops_per_pair": float(one["n_loops"].mean()),
           "pairs_restarted_frac": float((one["n_loops"] > 1).mean()), "valid_frac": float((one["valid"] == 1).mean()),
           "roofline": roofline_block(flops, {**det, "iterations_per_pair": float(one["iterations"].mean())}, kern_ms,
                                      peaks[64], "fp64_cuda_core", "plicp2_kernel<64> (short scans in a batch of 16 384 jobs or more; <128> for the log once) + jump_tables_kernel, restarts as a second batch", ex),
           "e2e": {"value": B / e2e_dt, "unit": UNIT, "steps": args.e2e_steps,
                   "h2d_bytes_per_step": int(raw_h.nbytes + 2 * B * 4), "d2h_bytes_per_step": int(B * 128),
                   "api": "zhsm_plicp_batch(host scans, host job lists, host records, ZHSM_FILTER_SCANS); pageable host "
                          "memory, as the reference's callers hold PMScan::r",
                   "device_vs_host_path_identical": same}}
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        work_h = raw_h.copy()
        t0 = time.perf_counter()
        po.median_filter_batch(work_h, cores)
        o = po.zhcsm_fmatch_batch(work_h, ri_h, si_h, lz.fi_min, lz.dfi, lz.min_range_cm, lz.max_range_cm, nthreads=cores)
        dt = time.perf_counter() - t0
        blk["cpu_baseline"] = {"value": B / dt, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"the whole log ({B} pairs), median filter + sm_icp restatement, {dt:.2f} s",
                               "parity_on_sample": parity_block(one, o)}
    return blk


def measure_c2(zh, eng, dev, stream, peaks, args):
    """BASELINE configs[1]: zhpsm polar scan matching odometry over data/fr079.log (pm_psm, and pm_icp beside it)."""
    import torch
    from laser_slam_b200 import engine as E
    from oracle import pyoracle as po
    from tests.support import load_log, make_laser as fixture_laser
    g = load_log("fr079")
    lz = fixture_laser(g)
    eng.set_laser(lz)
    raw_h = np.ascontiguousarray(g["scans"])
    S, n = raw_h.shape
    B, w, rep = S - 1, int(lz.scan_window), 8
    scans = torch.from_numpy(raw_h.copy()).to(dev)
    _, bad, seg = eng.preprocess(scans, stream=stream)
    ri = torch.arange(0, B, dtype=torch.int32, device=dev).repeat(rep)
    ai = ri + 1
    res = torch.zeros((B * rep, 128), dtype=torch.uint8, device=dev)
    ri_h, ai_h = np.arange(0, B, dtype=np.int32), np.arange(1, S, dtype=np.int32)
    out = {"workload": "zhpsm odometry over data/fr079.log (1463 consecutive pairs, 361 beams): pm_psm and pm_icp from zero poses",
           "unit": "jobs/s", "dtype": "f32", "beams": int(n), "jobs_per_step": int(B * rep),
           "value_note": f"the log's pairs x {rep} in one batch on preprocessed device-resident scans"}
    cores = os.cpu_count() or 1
    ref = None
    if not args.no_cpu and po.have_ref():
        ref = po.RefPolarMatch(g["laser_name"], fi_min=np.float32(g["fi_min"]), max_range_cm=int(g["max_range"]))
        pre = [ref.preprocess(s) for s in raw_h]
        Rr, Rb, Rs = (np.stack([p[k] for p in pre]) for k in range(3))
    for name, algo in (("pm_psm", E.ALGO_PSM), ("pm_icp", E.ALGO_PM_ICP)):
        ms = time_events(lambda: eng.psm(scans, bad, seg, ri, ai, algo=algo, results=res, stream=stream), reps=5, warm=2)
        kern_ms = eng.last_kernel_ms()
        r = _records(zh, res)
        it = r["iterations"].astype(np.float64)
        if algo == E.ALGO_PSM:
            # per iteration one projection (40 n); even iterations search the orientation (2 (2w+1) n),
            # odd ones estimate the translation (14 n); the loop counter starts at 1
            flops = float((it * 40.0 * n + np.floor(it / 2) * 2.0 * (2 * w + 1) * n + (it - np.floor(it / 2)) * 14.0 * n).sum())
            formula = "sum_jobs [ it*40n + floor(it/2)*2(2w+1)n + ceil(it/2)*14n ] (SURVEY 8d)"
        else:
            # pm_icp is not given a formula in SURVEY 8d; stated here: projection 40 n + windowed nearest neighbour
            # 5*2w per beam + trim / foot points / Procrustes sums 60 per beam, per iteration
            flops = float((it * (40.0 * n + 5.0 * 2 * w * n + 60.0 * n)).sum())
            formula = "sum_jobs it*(40n + 10wn + 60n) (window NN of ZHPolar_Match.cpp:1792-1810; stated in bench.py)"
        # e2e: raw host scans -> pm_preprocessScan -> matcher -> host records, all through the C ABI with host arrays
        def e2e_once():
            hs, hb, hg = eng.preprocess(raw_h.copy(), stream=stream)
            return eng.psm(hs, hb, hg, ri_h, ai_h, algo=algo, stream=stream)
        e2e_once()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            rh = e2e_once()
        e2e_dt = (time.perf_counter() - t0) / args.e2e_steps
        blk = {"value": B * rep / (ms * 1e-3), "ms_per_step": ms, "iterations_per_job": float(it.mean()),
               "failed_frac": float((r["valid"] == 0).mean()),
               "roofline": roofline_block(flops, {"formula": formula, "scan_window": w}, kern_ms, peaks[32], "fp32_cuda_core",
                                          f"psm_kernel ({name})", extra={"note": "latency-bound by design: the reference's sequential FP32 summation orders "
                                                                                  "are kept so that results are bit-identical (DESIGN 5)"}),
               "e2e": {"value": B / e2e_dt, "unit": "jobs/s", "steps": args.e2e_steps,
                       "h2d_bytes_per_step": int(raw_h.nbytes + S * n * 3 + 2 * B * 4), "d2h_bytes_per_step": int(raw_h.nbytes + S * n * 3 + B * 128),
                       "api": "zhsm_preprocess_batch(host) + zhsm_psm_batch(host arrays): the log's 1463 pairs per step",
                       "device_vs_host_path_identical": bool(np.array_equal(rh["x"], r[:B]["x"]))}}
        if ref is not None:
            t0 = time.perf_counter()
            poses, st = ref.match_batch(0 if algo == E.ALGO_PSM else 1, Rr, Rb, Rs, ri_h, ai_h, nthreads=cores)
            dt = time.perf_counter() - t0
            okj = st == 0
            blk["cpu_baseline"] = {"value": B / dt, "unit": "jobs/s", "cores": cores, "kind": "reference",
                                   "sample": f"the reference's own {name} (src/zhpsm compiled unmodified, oracle/_ref/libzhpsm_ref.so), "
                                             f"the whole log ({B} pairs) on preprocessed scans, {dt:.2f} s",
                                   "parity_on_sample": {"status_eq": float(((r[:B]["status"] != 0) == (st != 0)).mean()),
                                                        "pose_bit_identical": float((r[:B]["x"].astype(np.float32)[okj] == poses[okj]).all(1).mean())}}
        elif not args.no_cpu:
            blk["cpu_baseline"] = None
            blk["cpu_baseline_note"] = "oracle/_ref/libzhpsm_ref.so is not present on this box"
        out[name] = blk
    out["value"], out["metric"] = out["pm_psm"]["value"], "pm_psm_matches_per_sec"
    return out


def measure_c3(zh, eng, dev, stream, peaks, args):
    """BASELINE configs[2]: zhicp point-to-point ICP loop-closure candidates on data/mit-cscail.log, 64 initial guesses
    per pair (tests/loop_cases.py builds the pair list and the Philox guesses).  A step = the cloud pool
    and the (pair, guess) job list through zhsm_icp_batch + zhsm_best_of_starts."""
    import torch
    from oracle import pyoracle as po
    from tests.loop_cases import C3_G, build_c3
    from tests.support import load_log, make_laser as fixture_laser
    g = load_log("mit-cscail")
    eng.set_laser(fixture_laser(g))
    t0 = time.perf_counter()
    nodes, clouds, pose, flags, pairs, init = build_c3(
        g, lambda nd: eng.submap_cloud(nd["scans"], nd["traj"])[:2], eng.gate_candidates)
    build_s = time.perf_counter() - t0
    G, P = C3_G, len(pairs)
    pts_h, off, cnt = po.pack_clouds(clouds)
    ri_h, ci_h = np.repeat(pairs[:, 1], G).astype(np.int32), np.repeat(pairs[:, 0], G).astype(np.int32)
    jobs_h = np.ascontiguousarray(init.reshape(-1, 3))
    B = ri_h.size
    pts, ri, ci, jobs = (torch.from_numpy(a).to(dev) for a in (pts_h, ri_h, ci_h, jobs_h))
    res = torch.zeros((B, 128), dtype=torch.uint8, device=dev)
    best = [None]

    def step():
        eng.icp(pts, off, cnt, ri, ci, jobs, results=res, stream=stream)
        best[0] = eng.best_of_starts(res, P, G, stream=stream)
    ms = time_events(step, reps=5, warm=2)
    kern_ms = time_events(lambda: eng.icp(pts, off, cnt, ri, ci, jobs, results=res, stream=stream), reps=3, warm=0)
    r = _records(zh, res)
    rec = _records(zh, best[0][1])
    # F_iter = N2*(8 + 5*N1) + 12*pairs for exhaustive NN; the reference algorithm itself queries every
    # `corresponding_points_decimation`-th point (5) per iteration, so N2 is the decimated count here
    N1, N2 = cnt[ri_h].astype(np.float64), np.ceil(cnt[ci_h] / 5.0)
    it = r["iterations"].astype(np.float64)
    flops = float((it * (N2 * (8.0 + 5.0 * N1) + 12.0 * r["nvalid"])).sum())

    def e2e_once():
        rr = eng.icp(pts_h, off, cnt, ri_h, ci_h, jobs_h, stream=stream)       # host pool, host jobs, host records
        return eng.best_of_starts(rr, P, G, stream=stream)
    e2e_once()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        bh = e2e_once()
    e2e_dt = (time.perf_counter() - t0) / args.e2e_steps
    err = np.abs(rec["x"] - init[:, 0])
    acc = rec["goodness"] >= 0.8
    blk = {"workload": "zhicp loop-closure candidates on data/mit-cscail.log, 64 initial guesses per pair (SURVEY 8d C3: nodes of 15 "
                       "scans every 5 scans, consecutive node pairs + candidates within 2 m / bbox overlap >= 0.4, guess 0 = GT "
                       "relative pose, guesses 1..63 = guess 0 (+) U(+-0.5 m, +-0.5 m, +-15 deg), Philox seed 20261104)",
           "metric": "zhicp_matches_per_sec", "unit": "jobs/s", "dtype": "f32", "value": B / (ms * 1e-3), "ms_per_step": ms,
           "nodes": len(nodes), "pairs": int(P), "loop_candidates": int(P - (len(nodes) - 1)), "starts": G, "jobs_per_step": int(B),
           "cloud_points_mean": float(cnt.mean()), "iterations_per_job": float(it.mean()),
           "setup_s": {"submap_clouds_gating_guesses": build_s},
           "results": {"pairs_accepted_goodness_0.8": float(acc.mean()),
                       "median_err_vs_gt_of_accepted": [float(np.median(np.hypot(err[acc, 0], err[acc, 1]))), float(np.median(err[acc, 2]))]},
           "roofline": roofline_block(flops, {"formula": "sum_jobs it*(ceil(N2/5)*(8 + 5*N1) + 12*pairs) (SURVEY 8d, with the decimation the "
                                                         "reference's options prescribe)"}, kern_ms, peaks[32], "fp32_cuda_core", "icp_kernel"),
           "e2e": {"value": B / e2e_dt, "unit": "jobs/s", "steps": args.e2e_steps,
                   "h2d_bytes_per_step": int(pts_h.nbytes + 2 * B * 4 + B * 24 + 2 * off.size * 4), "d2h_bytes_per_step": int(B * 128 + P * 132),
                   "api": "zhsm_icp_batch(host cloud pool, host jobs, host records) + zhsm_best_of_starts(host)",
                   "device_vs_host_path_identical": bool(np.array_equal(bh[1]["x"], rec["x"]))}}
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        K = min(B, 4096)
        idx = np.linspace(0, B - 1, K).astype(np.int64)
        t0 = time.perf_counter()
        o = po.mrpt_icp_batch(pts_h, off, cnt, ri_h[idx], ci_h[idx], jobs_h[idx], nthreads=cores)
        dt = time.perf_counter() - t0
        dp = np.abs(r[idx]["x"] - o["pose"])
        blk["cpu_baseline"] = {"value": K / dt, "unit": "jobs/s", "cores": cores, "kind": "port",
                               "sample": f"{K} (pair, guess) jobs spread over the batch, MRPT CICP restatement, {dt:.2f} s",
                               "parity_on_sample": {"pose_within_1e-4m_1e-5rad": float(((dp[:, 0] <= 1e-4) & (dp[:, 1] <= 1e-4) & (dp[:, 2] <= 1e-5)).mean()),
                                                    "iterations_eq": float((r[idx]["iterations"] == o["nIterations"]).mean())}}
    return blk


C5_BEAMS, C5_STARTS, C5_SEED = 2048, 16, 20261105


def c5_laser(zh):
    n, PM_PI = C5_BEAMS, 3.141592654
    lz = zh.Laser()
    lz.n = n; lz.fi_min = np.float32(-PM_PI); lz.dfi = np.float32((360 * PM_PI / 180) / (n - 1.0))
    lz.min_range_cm = 10; lz.max_range_cm = MAX_RANGE_CM; lz.min_valid_points = n // 5; lz.scan_window = n // 10
    return lz


def measure_c5(zh, eng, dev, stream, world, rank, P, steps, warmup, args, peak64):
    """BASELINE configs[4]: multi-start PL-ICP for loop closure.  Per GPU P scan pairs (default 7812 =
    62 500 / 8) of 2048-beam 360-degree scans with a true offset up to 0.5 m / 20 degrees, 16 first guesses per pair =
    noisy odometry (+) {-0.3, +0.3}^2 m x {-10, -3.3, +3.3, +10} degrees.  A step = median filter -> one
    zhsm_plicp_batch over all (pair, start) jobs -> zhsm_best_of_starts (lowest mean error) on the device -> [N>1] NCCL
    all-gather of the best record per pair (pose + covariance, 128 bytes).  Sharded by pair, all starts of a pair on
    one GPU.  Returns the JSON-line fields (rank 0; None elsewhere)."""
    import torch
    import torch.distributed as dist
    from laser_slam_b200 import engine as E
    from laser_slam_b200 import shard
    n, G = C5_BEAMS, C5_STARTS
    lz = c5_laser(zh)
    eng.set_laser(lz)
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
        eng.plicp(work, ri, si, first_guess=guess, results=res, stream=stream, multistart=args.c5_multistart)
        if record:
            e1.record()
            kern_events.append((e0, e1))
        best, rec = eng.best_of_starts(res, P, G, criterion=E.BEST_MEAN_ERROR, stream=stream)
        out[0], out[1] = best, (shard.all_gather_records(rec, world * P) if world > 1 else rec)

    for _ in range(warmup):
        step()
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(dev.index) if rank == 0 else None
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        step(record=True)
    t1.record()
    barrier()
    ms = t0.elapsed_time(t1)
    launches = eng.launch_count() - launches0 + steps
    clk = clocks.stop() if clocks else None
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in kern_events]))
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    value = world * P * G * steps / (ms * 1e-3)

    # end to end with HOST buffers: scans, job lists and first guesses go up, the best record per pair comes back
    host_raw = torch.empty((2 * P, n), dtype=torch.float32, pin_memory=True)
    host_raw.copy_(raw)
    host_np = host_raw.numpy()
    si_h = ri_h + 1

    def e2e_step():
        r_dev = torch.empty((P * G, 128), dtype=torch.uint8, device=dev)
        eng.plicp(host_np, ri_h, si_h, first_guess=guess_h, results=r_dev, filter_scans=True, stream=stream, multistart=args.c5_multistart)
        _, rec = eng.best_of_starts(r_dev, P, G, criterion=E.BEST_MEAN_ERROR, stream=stream)
        if world > 1:
            rec = shard.all_gather_records(rec, world * P)
        return rec.cpu()

    e2e_step()
    barrier()
    w0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - w0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = world * P * G * args.e2e_steps / float(e2e_s.item())

    r = _records(zh, res).reshape(P, G)
    b = _records(zh, out[1])[rank * P:(rank + 1) * P]
    err = np.abs(b["x"] - tr); err[:, 2] = np.abs((err[:, 2] + np.pi) % (2 * np.pi) - np.pi)
    ok = (b["valid"] == 1) & (err[:, 0] < 0.02) & (err[:, 1] < 0.02) & (err[:, 2] < np.deg2rad(0.5))
    if rank != 0:
        return None
    sample = work[1: 2 * min(P, 256): 2].cpu().numpy()
    flops, fdet = plicp_flops(sample, n, lz.dfi, lz.min_range_cm, lz.max_range_cm, float(r["iterations"].sum()), P * G)
    ex = plicp_executed_flops(r.reshape(-1), fdet["n_s"], n)
    line = {"metric": "plicp_multistart_matches_per_sec", "value": value, "unit": "matches/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True,
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
            "roofline": roofline_block(flops, {**fdet, "iterations_per_start": float(r["iterations"].mean())}, kern_ms, peak64,
                                       "fp64_cuda_core", ("plicp2ms::plicp2_kernel<128> (4 starts of a pair per CTA)" if args.c5_multistart else "plicp2_kernel<256>") + " + jump_tables_kernel", ex),
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
        line["cpu_baseline"] = {"value": K / dt, "unit": "matches/s", "cores": cores, "kind": "port",
                                "sample": "the first %d (pair, start) jobs of the same batch, sm_icp restatement, %.2f s" % (K, dt),
                                "parity_on_sample": parity_block(r.reshape(-1)[:K], o)}
    return line


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
    # two record buffers: while step k+1's kernel fills one, step k's records are being all-gathered out of the other
    res2 = [torch.zeros((P, 128), dtype=torch.uint8, device=dev) for _ in range(2)]
    gath = shard.OverlappedGather(world * P, res2[0], depth=2)
    last = [0]
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    kern_events = []

    def step(record=False):
        k = gath.slot()
        work.copy_(raw)
        eng.median_filter(work, stream=stream)
        e0 = e1 = None
        if record:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        eng.plicp(work, ri, si, results=res2[k], stream=stream)
        if record:
            e1.record()
            kern_events.append((e0, e1))
        gath.submit(res2[k])
        last[0] = k

    for _ in range(args.warmup):
        step()
    gath.result(last[0])
    barrier()
    launches0 = eng.launch_count()
    clocks = ClockSampler(local) if rank == 0 else None
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start.record()
    for _ in range(args.steps):
        step(record=True)
    gathered = gath.result(last[0])     # the last step's gather is inside the timed region
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
    res = res2[last[0]]
    gather_ok = bool(torch.equal(gathered[rank * P:(rank + 1) * P], res)) if world > 1 else True

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
    eng2.close()

    if rank == 0:
        r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
        # parity of the device path vs the host-buffer path on the same inputs (bit-identical records)
        same = bool((r["corr_hash"] == res_np["corr_hash"]).all() and (r["iterations"] == res_np["iterations"]).all())
        it_sum = int(r["iterations"].sum())
        ns = min(P, 2048)
        flops, fl_detail = plicp_flops(host_np[1: 2 * ns: 2], n, lz.dfi, lz.min_range_cm, lz.max_range_cm, it_sum, P)
        ex = plicp_executed_flops(r, fl_detail["n_s"], n)
        peaks = {64: eng.fp_peak_tflops(64), 32: eng.fp_peak_tflops(32)}
        achieved = flops / (kern_ms * 1e-3) / 1e12
        hbm_bytes = P * (2 * n * 4 + 128)
        mp = {}
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(mp.get("hbm_gbs", 6650.0))
        roof = roofline_block(flops, {**fl_detail, "iterations_per_pair": it_sum / P}, kern_ms, peaks[64], "fp64_cuda_core",
                              "plicp2_kernel<192> + jump_tables_kernel", ex,
                              extra={"traffic": NCU_DRAM_BYTES_PER_PAIR * P,
                                     "traffic_source": f"ncu --set full dram__bytes_read+write of a 20k-pair launch ({NCU_DRAM_CAPTURE}), "
                                                       "scaled by pairs per launch",
                                     "fp32_peak": peaks[32], "frac_of_fp32_peak": achieved / peaks[32] if peaks[32] > 0 else None,
                                     "hbm": {"achieved_gbs": hbm_bytes / (kern_ms * 1e-3) / 1e9, "peak_gbs": hbm_peak,
                                             "frac": hbm_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak,
                                             "algorithmic_bytes_per_launch": hbm_bytes}})
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(P, world),
                "clocks": clk,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": args.e2e_steps,
                        "api": "zhsm_plicp_batch(host buffers, ZHSM_FILTER_SCANS), 2 handles on 2 host threads, chunks of %d pairs; "
                               "PINNED host memory (a caller holding pageable PMScan::r arrays pays the driver's staging copy on top)" % chunk,
                        "device_vs_host_path_identical": same},
                "gpu_launches": int(launches),
                "roofline": roof,
                "results": {"valid_frac": float((r["valid"] == 1).mean()),
                            "median_abs_err_vs_truth": np.median(np.abs(r["x"] - truth.cpu().numpy()), 0).tolist()}}
        if world > 1:
            line["all_gather"] = {"records_per_rank": P, "bytes_received_per_rank": int(world * P * 128),
                                  "overlap": "side stream, against the next step's kernel (double-buffered records)",
                                  "own_block_identical": gather_ok}
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            cp = min(P, args.cpu_pairs)
            v, dt, ores = cpu_time_sample(host_np[: 2 * cp], lz, cores)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"first {cp} pairs of the same batch, median filter + sm_icp restatement, {dt:.1f} s",
                                    "parity_on_sample": parity_block(r[:cp], ores)}
        if world == 1 and not args.no_configs:
            del raw, work, host_raw, host_np
            torch.cuda.empty_cache()
            cfgs = {}
            for name, fn in (("c1", measure_c1), ("c2", measure_c2), ("c3", measure_c3)):
                try:
                    cfgs[name] = fn(zh, eng, dev, stream, peaks, args)
                except Exception as e:  # the headline line must not depend on the side measurements
                    cfgs[name] = {"error": repr(e)}
            try:
                c5 = measure_c5(zh, eng, dev, stream, 1, 0, args.c5_pairs, 3, 3, args, peaks[64])
                c5.pop("clocks", None)
                cfgs["c5"] = c5
            except Exception as e:
                cfgs["c5"] = {"error": repr(e)}
            line["configs"] = cfgs
            eng.set_laser(lz)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_c5(args):
    """--workload c5: config 5 as a JSON line of its own (also under torchrun, sharded by pair)."""
    import torch
    import torch.distributed as dist
    import laser_slam_b200 as zh
    world, rank, local, dev = init_ranks()
    eng = zh.Engine(local)
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    P = args.pairs if args.pairs != 1_000_000 else args.c5_pairs
    eng.set_laser(c5_laser(zh))
    peak64 = eng.fp_peak_tflops(64)
    line = measure_c5(zh, eng, dev, tstream.cuda_stream, world, rank, P, args.steps, args.warmup, args, peak64)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=1_000_000, help="pairs per GPU per step (configs[3]: 1M)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-chunk", type=int, default=125_000)
    ap.add_argument("--cpu-pairs", type=int, default=100_000, help="cpu_baseline sample (pairs)")
    ap.add_argument("--ref-pairs", type=int, default=20_000, help="--impl reference: pairs per step")
    ap.add_argument("--c5-pairs", type=int, default=7812, help="config 5: pairs per GPU (62 500 / 8)")
    ap.add_argument("--c5-multistart", action="store_true",
                    help="config 5 through ZHSM_MULTISTART (four starts of a pair per CTA); measured SLOWER (68 k vs 90 k "
                         "matches/s, DESIGN 9), so off")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the C1/C2/C3/C5 block of the default line")
    ap.add_argument("--workload", choices=["c4", "c5"], default="c4",
                    help="c4 (default): the headline metric, 1080-beam odometry pairs (+ the configs block); c5: 2048-beam "
                         "multi-start loop closure as a line of its own")
    args = ap.parse_args()
    args.e2e_steps = max(1, args.e2e_steps)
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "c5":
        run_c5(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
