#!/bin/bash
# hd-256 backward with two roles per KV tile: parity of everything that touches hd 256, then C4 / C2 with and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -x -m gpu -k "256 or gemma or c4 or C4 or timed or flash or gpt" -p no:cacheprovider 2>&1 | tail -6
for wl in c4 c2; do
  python bench.py --workload $wl --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline --graph off > gpurun_out/r02aa_$wl.json 2> gpurun_out/r02aa_$wl.err || tail -5 gpurun_out/r02aa_$wl.err
  SPB_BWD256_ROLES=0 python bench.py --workload $wl --steps 20 --warmup 5 --no-e2e --sustain-s 0 --no-cpu-baseline --graph off > gpurun_out/r02aa_${wl}_cols.json 2>> gpurun_out/r02aa_$wl.err
done
python - <<'PY'
import json
for f in ("c4","c4_cols","c2","c2_cols"):
    try:
        d=json.loads(open(f"gpurun_out/r02aa_{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],4), "flash_bwd", d["stages_ms"].get("flash_bwd"))
    except Exception as e: print(f, "ERR", e)
PY
