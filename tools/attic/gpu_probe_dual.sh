#!/bin/bash
# Tuning probe: two interleaved walks per thread (plicp2_dual.cuh) in the 128-thread instance, against the same
# instance without them (variant d0) and against 192 threads, with and without chunk-order first iterations.
for n in 1080 840 361; do
  P=20000; [ $n -lt 500 ] && P=60000
  echo "$n beams, 192 threads: $(ZHSM_PLICP_T=192 PADS_KB=0 python tools/gpu_probe_occupancy.py $P $n | cut -c1-75)"
  for v in d0 ""; do for ci in 2 0; do
    echo "$n beams, 128 threads, variant=[$v] chunk_iters=$ci: $(ZHSM_LIB_VARIANT=$v ZHSM_PLICP_T=128 ZHSM_PLICP_CHUNK_ITERS=$ci PADS_KB=0 python tools/gpu_probe_occupancy.py $P $n | cut -c1-75)"
  done; done
done
