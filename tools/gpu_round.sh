#!/bin/bash
# One GPU-box visit: diagnostics + tests + bench, each in its own process with its own timeout; logs to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/smi.log 2>&1
timeout 300 python tools/diag_gemm.py > gpurun_out/diag_gemm.log 2>&1; echo "diag_gemm rc=$?" >> gpurun_out/diag_gemm.log
timeout 900 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "not tcgen05" --timeout 300 -p no:cacheprovider > gpurun_out/t1.log 2>&1
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "tcgen05" --timeout 120 -p no:cacheprovider > gpurun_out/t2.log 2>&1
timeout 1200 python -m pytest tests/test_gpu_block.py tests/test_gpu_dropin.py tests/test_gpu_dp.py -q -m gpu --timeout 600 -p no:cacheprovider > gpurun_out/t3.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err
for f in diag_gemm t1 t2 t3; do echo "=== $f"; tail -n 40 gpurun_out/$f.log; done
echo "=== bench"; cat gpurun_out/bench.log; tail -n 15 gpurun_out/bench.err
