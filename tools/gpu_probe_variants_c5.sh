#!/bin/bash
# Tuning probe: the C5 workload (and the 2048-beam good-guess batch) through tuning builds of the library.
for v in "" $@; do
  echo "variant=[$v] $(ZHSM_LIB_VARIANT=$v python tools/gpu_probe_c5.py 4096 2>&1 | grep -i "matches/s, \|digest" | cut -c1-150 | tr '\n' ' ')"
  echo "variant=[$v] 2048 beams, first guess 0: $(ZHSM_LIB_VARIANT=$v PADS_KB=0 python tools/gpu_probe_occupancy.py 8000 2048 | cut -c1-90)"
done
