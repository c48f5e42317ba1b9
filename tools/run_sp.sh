timeout 900 python -m pytest tests/test_gpu_blend_proof.py tests/test_gpu_strip.py tests/test_gpu_compose.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -8
STITCH_SINGLE_PASS=1 timeout 300 python tools/fused_prof.py c3 4
STITCH_SINGLE_PASS=1 timeout 300 python tools/fused_prof.py c3nb 4
STITCH_SINGLE_PASS=2 timeout 300 python tools/fused_prof.py c2 4
for w in c3 c3nb; do STITCH_SINGLE_PASS=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_sp_$w.csv python tools/fused_prof.py $w 2 > /dev/null 2>&1; done
STITCH_SINGLE_PASS=2 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_sp_c2.csv python tools/fused_prof.py c2 2 > /dev/null 2>&1
