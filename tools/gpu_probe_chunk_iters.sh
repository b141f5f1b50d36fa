#!/bin/bash
# Tuning probe: iterations searched in chunked order (thread t walks rays [t*C, (t+1)*C) and hands last_best on).
for ci in 0 1 2 3 4 6 50; do
  echo "chunk_iters=$ci C4: $(ZHSM_PLICP_CHUNK_ITERS=$ci PADS_KB=0 python tools/gpu_probe_occupancy.py 20000 1080 | cut -c1-75)"
  echo "chunk_iters=$ci C5: $(ZHSM_PLICP_CHUNK_ITERS=$ci python tools/gpu_probe_c5.py 2048 2>&1 | grep -i "matches/s\|walks per start" | cut -c1-170 | tr '\n' ' ')"
done
