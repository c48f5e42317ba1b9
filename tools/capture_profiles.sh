set -x
N="ncu --set full --clock-control none --import-source on"
timeout 600 $N -k regex:fused_kernel -c 1 -f -o gpurun_out/r2e_fused_c2 python tools/fused_prof.py c2 1 > gpurun_out/cap1.log 2>&1
timeout 600 $N -k regex:'compose_plan_kernel|fused_kernel' -c 2 -f -o gpurun_out/r2e_c3_blend python tools/fused_prof.py c3 1 > gpurun_out/cap2.log 2>&1
timeout 600 $N -k regex:'fused_plan_iv_kernel|fused_prep_kernel|fused_kernel' -c 3 -f -o gpurun_out/r2e_c3_noblend python tools/fused_prof.py c3nb 1 > gpurun_out/cap3.log 2>&1
timeout 600 $N -k regex:'jpeg_' -c 6 -f -o gpurun_out/r2e_jpeg python tools/jpeg_bench.py > gpurun_out/cap4.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2e_jpeg_launches.csv python tools/jpeg_bench.py > gpurun_out/cap5.log 2>&1
for f in gpurun_out/cap*.log; do tail -n 2 "$f"; done
