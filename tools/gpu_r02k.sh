#!/bin/bash
mkdir -p gpurun_out
for wl in c1 c2 c2lit c4 c4lit c5; do
  timeout 600 python bench.py --workload $wl --steps 20 --warmup 5 --sustain-s 0 > gpurun_out/r02k_bench_$wl.json 2> gpurun_out/r02k_bench_$wl.err; echo "bench $wl rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02k_bench_$wl.json"))
print("$wl ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), "TF/s", round(d["block_tflops"]["tflops"],1), "frac_burst", round(d["block_tflops"]["frac_of_burst_peak_per_gpu"],3), "e2e Mtok/s", round(d["e2e"]["value"]/1e6,3), "cpu", round(d["cpu_baseline"]["value"]))
print("   ", {k:v for k,v in d["stages_ms"].items() if v>0.004})
PY
done
