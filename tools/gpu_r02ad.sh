#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu -p no:cacheprovider 2>&1 | tail -2
for wl in c3 c5 c4 c2; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline --graph off > gpurun_out/r02ad_${wl}.json 2>> gpurun_out/r02ad.err
done
python - <<'PY'
import json
for wl in ("c3","c5","c4","c2"):
    try:
        d=json.loads(open(f"gpurun_out/r02ad_{wl}.json").read().strip().splitlines()[-1])
        print(wl, round(d["ms_per_step"],4), "flash_bwd", d["stages_ms"].get("flash_bwd"), "flash_fwd", d["stages_ms"].get("flash_fwd"))
    except Exception as e: print(wl, "ERR", e)
PY
tail -3 gpurun_out/r02ad.err
