for v in 0 1 2 3; do echo "variant $v"; STITCH_PREP_VARIANT=$v STITCH_SINGLE_PASS=1 timeout 300 python tools/fused_prof.py c3 4; STITCH_PREP_VARIANT=$v STITCH_SINGLE_PASS=1 timeout 300 python tools/fused_prof.py c3nb 4; done
STITCH_PREP_VARIANT=3 STITCH_SINGLE_PASS=2 timeout 300 python -m pytest tests/test_gpu_strip.py -x -q 2>&1 | tail -2
