#!/bin/bash
# tools/profile_round.sh <tag> -- run on the GPU box (under gpurun): the profiling recipe of B200_PROFILING.md for the
# three matcher kernels.  Every program is first run plainly (must exit 0), then once under ncu.  Outputs go to
# gpurun_out/<tag>_*; the summaries that are judged are copied into profiles/ by hand afterwards.
set -u
tag=${1:-rXX}
out=gpurun_out
mkdir -p $out
BENCH="python bench.py --steps 2 --warmup 1 --pairs 20000 --no-cpu --e2e-steps 1"
$BENCH > $out/${tag}_plain.log 2>&1 || { echo "plain bench failed"; tail -n 5 $out/${tag}_plain.log; exit 1; }
# launch list (cold-cache, serialised per-launch durations): the kernel's SHARE of the step is what must agree
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $BENCH > $out/${tag}_ncu1.log 2>&1
# one full capture of the dominant kernel
ncu --set full --clock-control none --import-source on -k regex:plicp2_kernel -c 1 -f -o $out/${tag}_plicp $BENCH > $out/${tag}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:jump_tables_kernel -c 1 -f -o $out/${tag}_jump $BENCH > $out/${tag}_ncu2b.log 2>&1
# the zhpsm and zhicp kernels: plain, then one full capture each
python tools/gpu_probe_psm.py fr079 4 > $out/${tag}_psm_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:psm_kernel -c 2 -f -o $out/${tag}_psm python tools/gpu_probe_psm.py fr079 4 > $out/${tag}_ncu3.log 2>&1
python tools/gpu_probe_icp.py 2 > $out/${tag}_icp_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:icp_kernel -c 1 -f -o $out/${tag}_icp python tools/gpu_probe_icp.py 2 > $out/${tag}_ncu4.log 2>&1
ls -la $out | grep ${tag}_
