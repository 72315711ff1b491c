#!/bin/bash
# bf16 decode: mma.sync weight-stream GEMM + mma.sync attention; parity then benches and a launch list
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_decode.py tests/test_gpu_dropin.py -q -x 2>&1 | tail -6
for wl in c3dec c3dec32; do
  python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02v_$wl.json 2> gpurun_out/r02v_$wl.err || tail -5 gpurun_out/r02v_$wl.err
  SPB_DECODE_GEMM_MMA=0 python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02v_${wl}_tc.json 2>> gpurun_out/r02v_$wl.err
done
python - <<'PY'
import json
for f in ("c3dec","c3dec_tc","c3dec32","c3dec32_tc"):
    try:
        d=json.loads(open(f"gpurun_out/r02v_{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],4), round(d["roofline"]["frac"],3), round(d["e2e"]["ms_per_step"],4))
    except Exception as e: print(f, "ERR", e)
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r02v_ncu_launches_c3dec.csv python bench.py --workload c3dec --steps 4 --warmup 3 > gpurun_out/r02v_ncu.log 2>&1
tail -14 gpurun_out/r02v_ncu_launches_c3dec.csv | cut -d, -f5,12- | cut -c1-200
