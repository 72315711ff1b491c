#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_timed_shapes.py tests/test_gpu_block.py -q --timeout 900 -p no:cacheprovider -x -k "flash or block or gpt or c3" > gpurun_out/r02l_tests.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/r02l_tests.log
python tools/diag_fb_perf.py 2>&1 | tee gpurun_out/r02l_fb_perf.log
