#!/bin/bash
# A/B of library builds on everything: uniform batches at 1080 / 360 / 1440 beams, the C5 probe, the three real logs.
for v in "" $@; do
  echo "== variant=[$v]"
  ZHSM_LIB_VARIANT=$v tools/gpu_ab.sh 2>&1 | cut -c1-170
  ZHSM_LIB_VARIANT=$v python tools/gpu_probe_logs.py 2>&1 | tail -3 | cut -c1-200
done
