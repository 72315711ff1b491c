#!/bin/bash
# decode chain with programmatic dependent launch: parity (repeated, races would show as flaky values), benches with / without
mkdir -p gpurun_out
for i in 1 2 3; do timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_dropin.py -q -x 2>&1 | tail -2; done
for wl in c3dec c3dec32 c1dec; do
  python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02w_$wl.json 2> gpurun_out/r02w_$wl.err || tail -5 gpurun_out/r02w_$wl.err
  SPB_DECODE_PDL=0 python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02w_${wl}_nopdl.json 2>> gpurun_out/r02w_$wl.err
done
python - <<'PY'
import json
for f in ("c3dec","c3dec_nopdl","c3dec32","c3dec32_nopdl","c1dec","c1dec_nopdl"):
    try:
        d=json.loads(open(f"gpurun_out/r02w_{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["e2e"]["ms_per_step"],4))
    except Exception as e: print(f, "ERR", e)
PY
