#!/bin/bash
# multi-GPU sanity: the 2-rank gradient test + bench at N (default path)
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dp.py -q --timeout 600 -p no:cacheprovider > gpurun_out/r02n_tests_dp.log 2>&1; echo "dp pytest rc=$?"; tail -3 gpurun_out/r02n_tests_dp.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02n_bench_c3_${N}gpu.json 2> gpurun_out/r02n_bench_${N}gpu.err; echo "bench N=$N rc=$?"
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r02n_bench_c3_${N}gpu.json") if l.startswith("{")][-1])
print("N=$N ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), "e2e Mtok/s", round(d["e2e"]["value"]/1e6,3), "host ceiling", d["e2e"].get("host_ceiling_gbs",{}).get("h2d"), d["e2e"].get("host_ceiling_gbs",{}).get("d2h"), "grads", d["config"]["grad_buffers"], "dp_check", d["dp_check"]["max_rel_err"], "sustained", (d.get("sustained") or {}).get("ms_per_step"))
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r02n_reference_${N}gpu.json 2>> gpurun_out/r02n_bench_${N}gpu.err; echo "reference N=$N rc=$?"; tail -c 300 gpurun_out/r02n_reference_${N}gpu.json
