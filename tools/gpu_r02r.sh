#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -q -m gpu --timeout 900 -p no:cacheprovider -x > gpurun_out/r02r_tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/r02r_tests_gpu.log
for wl in c1 c2 c5; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --sustain-s 0 --no-cpu-baseline > gpurun_out/r02r_bench_$wl.json 2> gpurun_out/r02r_bench_$wl.err; echo "bench $wl rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02r_bench_$wl.json"))
print("$wl ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), {k:v for k,v in d["stages_ms"].items() if v>0.004})
PY
done
