#!/bin/bash
mkdir -p gpurun_out
python bench.py --workload c3 --graph on --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline > gpurun_out/r02t_c3.json 2> gpurun_out/r02t.err; tail -3 gpurun_out/r02t.err
python bench.py --workload c2 --graph on --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline > gpurun_out/r02t_c2.json 2>> gpurun_out/r02t.err
python bench.py --workload c4 --graph on --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline > gpurun_out/r02t_c4.json 2>> gpurun_out/r02t.err
python - <<'PY'
import json
for f in ("c3","c2","c4"):
    try:
        d=json.loads(open(f"gpurun_out/r02t_{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d.get("graph"), d["clocks"].get("sm_mhz"))
    except Exception as e: print(f, "ERR", e)
PY
