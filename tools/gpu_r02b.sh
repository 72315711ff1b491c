#!/bin/bash
# round 2, visit B: decode tests + decode bench lines (+ full GPU suite)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_nll.py -q --timeout 600 -p no:cacheprovider -x > gpurun_out/r02b_tests_decode.log 2>&1; echo "decode pytest rc=$?"
tail -n 25 gpurun_out/r02b_tests_decode.log
for wl in c3dec c3dec32 c1dec; do
  timeout 600 python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02b_bench_$wl.json 2> gpurun_out/r02b_bench_$wl.err; echo "bench $wl rc=$?"
  cat gpurun_out/r02b_bench_$wl.json; tail -n 3 gpurun_out/r02b_bench_$wl.err
done
timeout 1500 python -m pytest tests -q -m gpu --timeout 900 -p no:cacheprovider > gpurun_out/r02b_tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/r02b_tests_gpu.log
