#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_block.py tests/test_gpu_dropin.py tests/test_gpu_dp.py tests/test_gpu_timed_shapes.py -q --timeout 900 -p no:cacheprovider -k "not gemm_pair and not flash_at" > gpurun_out/r02p_tests.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/r02p_tests.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
SPB_F32_TC=0 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
for wl in c1 c2lit; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --sustain-s 0 --no-cpu-baseline > gpurun_out/r02p_bench_$wl.json 2> gpurun_out/r02p_bench_$wl.err; echo "bench $wl rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02p_bench_$wl.json"))
print("$wl ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), {k:v for k,v in d["stages_ms"].items() if v>0.004})
PY
done
