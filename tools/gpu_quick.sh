#!/bin/bash
# quick GPU visit: full gpu test-suite, flash diag (timings), bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -p no:cacheprovider -x > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/tests_gpu.log
timeout 300 python tools/diag_flash.py --bwd > gpurun_out/diag_flash.log 2>&1; echo "diag rc=$?"; grep -E "time|FAIL|EXC|ALL" gpurun_out/diag_flash.log
timeout 900 python bench.py --steps 20 --warmup 5 "$@" > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.log; tail -n 5 gpurun_out/bench.err
