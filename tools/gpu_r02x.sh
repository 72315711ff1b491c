#!/bin/bash
# recorded decode step (spb_block_decode_dev in a CUDA graph): parity, then benches
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode.py tests/test_gpu_dropin.py -q -x 2>&1 | tail -6
for wl in c3dec c3dec32 c1dec; do
  python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02x_$wl.json 2> gpurun_out/r02x_$wl.err || tail -5 gpurun_out/r02x_$wl.err
done
python - <<'PY'
import json
for f in ("c3dec","c3dec32","c1dec"):
    try:
        d=json.loads(open(f"gpurun_out/r02x_{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["e2e"]["ms_per_step"],4), d.get("graph"))
    except Exception as e: print(f, "ERR", e)
PY
