#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_block.py tests/test_gpu_timed_shapes.py tests/test_gpu_dropin.py tests/test_gpu_ffn.py -q --timeout 900 -p no:cacheprovider -x -k "not gemm_pair and not flash_at" > gpurun_out/r02f_tests.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/r02f_tests.log
for f in 1 0 1 0; do
  SPB_BWD_FUSE_PREP=$f timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --sustain-s 0 > gpurun_out/r02f_bench_fuse$f.json 2> gpurun_out/r02f_bench.err; echo "bench fuse=$f rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02f_bench_fuse$f.json"))
s=d["stages_ms"]
print("fuse=$f ms/step", round(d["ms_per_step"],4), "dattn", s["gemm_dattn"], "flash_bwd", s["flash_bwd"], "dwo", s["gemm_dwo"], "sum", round(s["gemm_dattn"]+s["flash_bwd"],4))
PY
done
