#!/bin/bash
# round 2, visit A: full GPU test suite, C3 bench line (with sustained leg + host ceiling), compute-sanitizer passes on small shapes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/r02a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout 900 -p no:cacheprovider -x > gpurun_out/r02a_tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/r02a_tests_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02a_bench_c3.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"
cat gpurun_out/r02a_bench_c3.json; tail -n 5 gpurun_out/r02a_bench.err
SEL='test_attention_tcgen05_flash or test_gemm_tcgen05_bf16 or test_gemm_tcgen05_epilogues or test_gemm_tcgen05_split_k'
for tool in memcheck synccheck racecheck; do
  timeout 420 compute-sanitizer --tool $tool --log-file gpurun_out/r02a_sanitizer_$tool.log python -m pytest tests/test_gpu_kernels.py -q -x -p no:cacheprovider -k "$SEL" > gpurun_out/r02a_sanitizer_${tool}_pytest.log 2>&1
  echo "sanitizer $tool rc=$?"; tail -n 3 gpurun_out/r02a_sanitizer_${tool}_pytest.log; tail -n 4 gpurun_out/r02a_sanitizer_$tool.log
done
