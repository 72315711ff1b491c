#!/bin/bash
# bf16 decode attention on mma.sync: parity, then the decode benches (old kernel via SPB_DECODE_MMA=0 for comparison)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_decode.py -q -x 2>&1 | tail -8
for wl in c3dec c3dec32; do
  python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02u_$wl.json 2> gpurun_out/r02u_$wl.err || tail -5 gpurun_out/r02u_$wl.err
  SPB_DECODE_MMA=0 python bench.py --workload $wl --steps 50 --warmup 5 > gpurun_out/r02u_${wl}_simt.json 2>> gpurun_out/r02u_$wl.err
done
python - <<'PY'
import json
for f in ("c3dec","c3dec_simt","c3dec32","c3dec32_simt"):
    try:
        d=json.loads(open(f"gpurun_out/r02u_{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],4), d["roofline"]["frac"], d["roofline"].get("attn_decode_kernel"))
    except Exception as e: print(f, "ERR", e)
PY
