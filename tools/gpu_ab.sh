#!/bin/bash
# A/B of library builds ("" = the product, then the variants named on the command line): C4 kernel time (20 k pairs,
# 1080 beams) + digest, other beam counts, the C5 probe.
for v in "" $@; do
  echo "variant=[$v] C4: $(ZHSM_LIB_VARIANT=$v PADS_KB=0 python tools/gpu_probe_occupancy.py 20000 1080 | cut -c1-80)"
  echo "variant=[$v] 360 beams: $(ZHSM_LIB_VARIANT=$v PADS_KB=0 python tools/gpu_probe_occupancy.py 40000 360 | cut -c1-80)"
  echo "variant=[$v] 1440 beams: $(ZHSM_LIB_VARIANT=$v PADS_KB=0 python tools/gpu_probe_occupancy.py 10000 1440 | cut -c1-80)"
  echo "variant=[$v] $(ZHSM_LIB_VARIANT=$v python tools/gpu_probe_c5.py 4096 2>&1 | grep -i "matches/s, \|digest" | cut -c1-150 | tr '\n' ' ')"
done
