#!/bin/bash
# GEMM epilogue through shared memory + TMA stores: full parity, then the configs with and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu -p no:cacheprovider 2>&1 | tail -2
python tools/diag_gemm_bias.py
for wl in c3 c5 c2 c4 c1; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline --graph off > gpurun_out/r02af_${wl}.json 2>> gpurun_out/r02af.err
  SPB_GEMM_TMA_STORE=0 timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline --graph off > gpurun_out/r02af_${wl}_direct.json 2>> gpurun_out/r02af.err
done
python - <<'PY'
import json
for wl in ("c3","c5","c2","c4","c1"):
  for v in ("","_direct"):
    try:
        d=json.loads(open(f"gpurun_out/r02af_{wl}{v}.json").read().strip().splitlines()[-1])
        g={k:round(x,4) for k,x in d["stages_ms"].items() if k.startswith("gemm")}
        print(wl+v, round(d["ms_per_step"],4), g)
    except Exception as e: print(wl+v, "ERR", e)
PY
tail -3 gpurun_out/r02af.err
