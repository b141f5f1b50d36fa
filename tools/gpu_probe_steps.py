import os, sys
import numpy as np, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import laser_slam_b200 as zh
import c5_workload
eng, scans, ri, si, guess = c5_workload.make(zh, 512)
res = torch.zeros((ri.numel(), 128), dtype=torch.uint8, device="cuda:0")
for _ in range(2):
    eng.plicp(scans, ri, si, first_guess=guess, results=res); eng.sync()
r = res.cpu().numpy().view(zh.RESULT_DTYPE).reshape(-1)
print("kernel ms", eng.last_kernel_ms(), "walks", r["reserved"][:,0].mean(), "first", r["reserved"][:,1].mean(), "rounds", r["reserved"][:,2].mean(), "steps", r["reserved"][:,3].mean())
