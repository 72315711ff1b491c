#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -q --timeout 600 -p no:cacheprovider -x > gpurun_out/r02d_tests_decode.log 2>&1; echo "decode pytest rc=$?"
tail -n 4 gpurun_out/r02d_tests_decode.log
CMD="python bench.py --workload c3dec --steps 3 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:attn_decode_kernel|small_m_gemm_kernel" -s 9 -c 3 -o gpurun_out/r02d_prof_decode -f $CMD > gpurun_out/r02d_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
