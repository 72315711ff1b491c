#!/bin/bash
# One GPU-box visit: tests + bench (+ optional ncu passes), each in its own process with its own timeout; logs to gpurun_out/.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 -p no:cacheprovider > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/tests_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.log; tail -n 5 gpurun_out/bench.err
if [ "$1" == "ncu" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  echo "ncu launches rc=$?"
  # warm-up 3 steps -> skip 3 steps' worth of each kernel family, then take one full step of it
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_tc2_kernel|gemm_tc_kernel" -s 18 -c 6 -o gpurun_out/prof_gemm -f $CMD > gpurun_out/ncu_gemm.log 2>&1
  echo "ncu gemm rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:flash_ -s 12 -c 4 -o gpurun_out/prof_flash -f $CMD > gpurun_out/ncu_flash.log 2>&1
  echo "ncu flash rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"norm_|colsum_|rope_" -s 12 -c 4 -o gpurun_out/prof_norm -f $CMD > gpurun_out/ncu_norm.log 2>&1
  echo "ncu norm rc=$?"
  ls -la gpurun_out/*.ncu-rep
fi
