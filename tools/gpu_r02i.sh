#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_block.py -q --timeout 600 -p no:cacheprovider -x -k "wide or gemma or 768" > gpurun_out/r02i_tests.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/r02i_tests.log
timeout 600 python bench.py --workload c4lit --steps 20 --warmup 5 --no-cpu-baseline --sustain-s 0 > gpurun_out/r02i_bench_c4lit.json 2> gpurun_out/r02i_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02i_bench_c4lit.json"))
print("c4lit ms/step", round(d["ms_per_step"],4), "tok/s", round(d["value"]), "block TF/s", round(d["block_tflops"]["tflops"],1), d["stages_ms"])
PY
tail -3 gpurun_out/r02i_bench.err
