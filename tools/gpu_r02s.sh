#!/bin/bash
# graph leg: captured-step test, C1 fp32 / bf16 and C5 with the replay leg
mkdir -p gpurun_out
python -m pytest tests/test_gpu_block.py -q -x -k captured 2>&1 | tail -5
python bench.py --workload c1 --steps 50 --warmup 5 --no-e2e --sustain-s 0 > gpurun_out/r02s_c1.json 2> gpurun_out/r02s_c1.err; tail -3 gpurun_out/r02s_c1.err
python bench.py --workload c5 --graph on --steps 20 --warmup 3 --no-e2e --sustain-s 0 > gpurun_out/r02s_c5.json 2>> gpurun_out/r02s_c1.err
python - <<'PY'
import json
for f in ("c1","c5"):
    try:
        d=json.loads(open(f"gpurun_out/r02s_{f}.json").read().strip().splitlines()[-1])
        print(f, d["ms_per_step"], d.get("graph"), d["config"].get("l2"), d["gpu_launches"])
    except Exception as e: print(f, "ERR", e)
PY
