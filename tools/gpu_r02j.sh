#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_nll.py -q --timeout 600 -p no:cacheprovider -x > gpurun_out/r02j_tests.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/r02j_tests.log
timeout 300 python tools/diag_nll.py > gpurun_out/r02j_diag_nll.log 2>&1; echo "diag rc=$?"; tail -12 gpurun_out/r02j_diag_nll.log
