#!/bin/bash
# A/B of the product build: C4 kernel time (20 k pairs, 1080 beams) + digest, C5 probe, intel-lab parity test.
echo "C4: $(PADS_KB=0 python tools/gpu_probe_occupancy.py 20000 1080)"
echo "C4/1440: $(PADS_KB=0 python tools/gpu_probe_occupancy.py 10000 1440)"
python tools/gpu_probe_c5.py 2>&1 | tail -4
