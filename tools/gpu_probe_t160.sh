#!/bin/bash
# Tuning probe: warps per SM.  The 160-thread instance (5 CTAs/SM where the footprint allows) and a 64-register
# build of the 192-thread one (variant n5: 5 CTAs/SM = 30 warps) against the product's 192 threads x 80 registers x 4.
for n in 840 1080; do
  for T in 192 160; do
    echo "n=$n T=$T: $(ZHSM_PLICP_T=$T PADS_KB=0 python tools/gpu_probe_occupancy.py 20000 $n)"
  done
  echo "n=$n T=192 64 regs: $(ZHSM_LIB_VARIANT=n5 ZHSM_PLICP_T=192 PADS_KB=0 python tools/gpu_probe_occupancy.py 20000 $n)"
done
