#!/bin/bash
mkdir -p gpurun_out
BASE=${BASE:-solvingpapers_b200/_ab/libspb200_base.so}
timeout 900 python -m pytest tests -q -m gpu --timeout 600 -p no:cacheprovider -x > gpurun_out/tests_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 8 gpurun_out/tests_gpu.log
run() {
  local name=$1; shift
  env "$@" timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_$name.log 2> gpurun_out/bench_$name.err; echo "bench $name rc=$?"
  python - "$name" <<'PY'
import json,sys
name=sys.argv[1]
for l in open(f'gpurun_out/bench_{name}.log'):
    if l.startswith('{'):
        d=json.loads(l); print(name, 'ms', round(d['ms_per_step'],4), 'tok/s', round(d['value']), 'e2e', round(d['e2e']['value']), {k:round(v,4) for k,v in d['stages_ms'].items()})
PY
}
if [ -f "$BASE" ]; then run base SPB_LIB_PATH=$PWD/$BASE; fi
run new A=1
if [ -f "$BASE" ]; then run base2 SPB_LIB_PATH=$PWD/$BASE; fi
run new2 A=1
timeout 300 python tools/diag_flash.py --bwd > gpurun_out/diag_flash.log 2>&1; echo "diag rc=$?"; grep -E "time|FAIL|EXC|ALL" gpurun_out/diag_flash.log
cp gpurun_out/bench_new2.log gpurun_out/bench.log
