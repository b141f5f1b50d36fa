"""Scratch probe: zhsm_preprocess_batch (median filter + far points + segmentation) on device-resident scans."""
import sys, os, torch, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import laser_slam_b200 as zh
from tests.support import load_log, make_laser
g = load_log("fr079"); eng = zh.Engine(0); eng.set_laser(make_laser(g))
dev = torch.device("cuda:0")
rep = 64
scans = torch.from_numpy(np.tile(g["scans"], (rep, 1))).to(dev)
S, n = scans.shape
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
for _ in range(2):
    x = scans.clone(); bad = torch.zeros((S, n), dtype=torch.uint8, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); _, b, sg = eng.preprocess(x, bad, stream=st.cuda_stream); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print(f"preprocess {S} x {n}: {ms:.3f} ms -> {S/ms*1e3:.0f} scans/s, {S*n*(8+2+2)/ms/1e6:.0f} GB/s of algorithmic traffic")
e0.record(); o = eng.is_corridor(x, b, sg, stream=st.cuda_stream); e1.record(); torch.cuda.synchronize()
print(f"is_corridor {S} x {n}: {e0.elapsed_time(e1):.3f} ms")
