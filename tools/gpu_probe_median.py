import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
import laser_slam_b200 as zh, bench
eng = zh.Engine(0); lz = bench.make_laser(zh); eng.set_laser(lz)
dev = torch.device('cuda:0')
for S in (40000, 2000000):
    x = torch.rand((S, lz.n), device=dev) * 3000
    st = torch.cuda.Stream(); torch.cuda.set_stream(st)
    eng.median_filter(x, stream=st.cuda_stream); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); eng.median_filter(x, stream=st.cuda_stream); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"median filter {S} x {lz.n}: {ms:.3f} ms -> {S*lz.n*8/ms/1e6:.0f} GB/s")
