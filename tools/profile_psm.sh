set -u
out=gpurun_out
python tools/gpu_probe_psm.py fr079 4 > $out/r01f_psm_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:psm_kernel -s 2 -c 1 -f -o $out/r01f_pm_icp python tools/gpu_probe_psm.py fr079 4 > $out/r01f_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:psm_kernel -s 0 -c 1 -f -o $out/r01f_pm_psm python tools/gpu_probe_psm.py fr079 4 > $out/r01f_ncu2.log 2>&1
ls -la $out | grep r01f
