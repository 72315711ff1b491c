#!/bin/bash
# round 2: the record visit - full GPU suite, headline bench (+ reference arm), ncu launch list and full captures of the top kernels
mkdir -p gpurun_out
T=r02z
timeout 1800 python -m pytest tests -q -m gpu --timeout 900 -p no:cacheprovider > gpurun_out/${T}_tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/${T}_tests_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_c3.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2>> gpurun_out/${T}_bench.err; echo "reference rc=$?"
python tools/show_bench.py gpurun_out/${T}_bench_c3.json 2>/dev/null | head -30
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --sustain-s 0 --no-host-ceiling"
timeout 600 $CMD > gpurun_out/${T}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_ncu_launches_c3.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm_tc2_kernel|gemm_tc_kernel" -s 18 -c 6 -o gpurun_out/${T}_prof_gemm -f $CMD > gpurun_out/${T}_ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:flash_ -s 9 -c 3 -o gpurun_out/${T}_prof_flash -f $CMD > gpurun_out/${T}_ncu_flash.log 2>&1; echo "ncu flash rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"norm_|colsum_|rope_" -s 12 -c 4 -o gpurun_out/${T}_prof_norm -f $CMD > gpurun_out/${T}_ncu_norm.log 2>&1; echo "ncu norm rc=$?"
ls -la gpurun_out/${T}_*.ncu-rep
