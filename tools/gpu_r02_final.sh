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
# decode: plain run, then the full capture of its two streaming kernels
DCMD="python bench.py --workload c3dec32 --steps 4 --warmup 3 --graph off --no-cpu-baseline"
timeout 600 $DCMD > gpurun_out/${T}_plain_dec.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:attn_decode_mma_kernel|small_m_mma_kernel" -s 9 -c 3 -o gpurun_out/${T}_prof_decode -f $DCMD > gpurun_out/${T}_ncu_decode.log 2>&1; echo "ncu decode rc=$?"
# hd-256 backward: the legacy column split still passes; full capture of the role kernel at C4
SPB_BWD256_ROLES=0 timeout 600 python -m pytest tests -q -m gpu -k "256 or c4 or C4 or gemma" -p no:cacheprovider 2>&1 | tail -1
C4CMD="python bench.py --workload c4 --steps 2 --warmup 3 --graph off --no-cpu-baseline --no-e2e --sustain-s 0"
timeout 600 $C4CMD > gpurun_out/${T}_plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:flash_bwd256_roles_kernel" -s 3 -c 1 -o gpurun_out/${T}_prof_bwd256 -f $C4CMD > gpurun_out/${T}_ncu_bwd256.log 2>&1; echo "ncu bwd256 rc=$?"
# the other configs and the decode lines on the final library, and the per-shape cuBLAS comparison
for wl in c1 c2 c2lit c4 c4lit c5 c3dec c3dec32 c1dec; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --sustain-s 0 > gpurun_out/${T}_bench_$wl.json 2>> gpurun_out/${T}_bench.err; echo "$wl rc=$?"
done
timeout 300 python tools/diag_gemm_vs_cublas.py > gpurun_out/${T}_gemm_vs_cublas.jsonl 2>> gpurun_out/${T}_bench.err; tail -1 gpurun_out/${T}_gemm_vs_cublas.jsonl
