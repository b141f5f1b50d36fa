#!/bin/bash
# tools/profile_round.sh <tag> <part> -- run on the GPU box (under gpurun): the profiling recipe of B200_PROFILING.md for
# every kernel family of the library, in two parts (a: PL-ICP, b: zhpsm / zhicp / graph) because gpurun brings back at
# most 64 MiB of gpurun_out/ per call.  Every program is first run plainly (must exit 0), then once under ncu.  Outputs go to
# gpurun_out/<tag>_*; the summaries that are judged are condensed into profiles/ afterwards (tools/ncu_summary.py,
# tools/ncu_hotspots.py).
set -u
tag=${1:-rXX}
part=${2:-a}
out=gpurun_out
mkdir -p $out
NCU="ncu --set full --clock-control none --import-source on"
BENCH="python bench.py --steps 2 --warmup 1 --pairs 20000 --no-cpu --no-configs --e2e-steps 1"
if [ "$part" = "a" ]; then
$BENCH > $out/${tag}_plain.log 2>&1 || { echo "plain bench failed"; tail -n 5 $out/${tag}_plain.log; exit 1; }
# launch list (cold-cache, serialised per-launch durations): the kernel's SHARE of the step is what must agree
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $BENCH > $out/${tag}_ncu1.log 2>&1
# one full capture of the dominant kernel (C4: 1080 beams) and of its jump-table kernel
$NCU -k regex:plicp2_kernel -c 1 -f -o $out/${tag}_plicp $BENCH > $out/${tag}_ncu2.log 2>&1
$NCU -k regex:jump_tables_kernel -c 1 -f -o $out/${tag}_jump $BENCH > $out/${tag}_ncu2b.log 2>&1
# C5: the 256-thread instance on 2048-beam multi-start jobs (512 pairs x 16 starts)
python tools/gpu_probe_c5.py 512 > $out/${tag}_c5_plain.log 2>&1 && \
  $NCU -k regex:plicp2_kernel -c 1 -f -o $out/${tag}_c5_plicp python tools/gpu_probe_c5.py 512 > $out/${tag}_ncu2c.log 2>&1
fi
if [ "$part" = "b" ]; then
# the zhpsm and zhicp kernels: plain, then one full capture each (pm_psm = first psm_kernel launch, pm_icp = third)
python tools/gpu_probe_psm.py fr079 4 > $out/${tag}_psm_plain.log 2>&1 && \
  $NCU -k regex:psm_kernel -c 3 -f -o $out/${tag}_psm python tools/gpu_probe_psm.py fr079 4 > $out/${tag}_ncu3.log 2>&1
python tools/gpu_probe_icp.py 2 > $out/${tag}_icp_plain.log 2>&1 && \
  $NCU -k regex:icp_kernel -c 1 -f -o $out/${tag}_icp python tools/gpu_probe_icp.py 2 > $out/${tag}_ncu4.log 2>&1
# the kernels either side of the ICP batches: submap clouds (raycast), gating, consensus
python tools/gpu_probe_graph.py mit-cscail > $out/${tag}_graph_plain.log 2>&1 && \
  $NCU -k regex:'raycast_kernel|gate_kernel|consensus_kernel|emit_kernel' -c 8 -f -o $out/${tag}_graph python tools/gpu_probe_graph.py mit-cscail > $out/${tag}_ncu5.log 2>&1
fi
ls -la $out | grep ${tag}_; du -sm $out
