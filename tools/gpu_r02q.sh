#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
run() {
  name=$1; shift
  envs=(); while [ "$1" != "--" ] && [ $# -gt 0 ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --sustain-s 0 "$@" > gpurun_out/r02q_n${N}_$name.json 2> gpurun_out/r02q_n${N}_$name.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r02q_n${N}_$name.json") if l.startswith("{")][-1])
    s=d["stages_ms"]
    print("  N=$N $name: ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), "| dx", s["gemm_dx"], "allreduce", s["allreduce"], "flash_bwd", s["flash_bwd"], "dwqkv", s["gemm_dwqkv"], "norm_bwd", s["norm_bwd"], "| dp_check", d.get("dp_check",{}).get("max_rel_err"))
except Exception as ex:
    print("  $name parse failed:", ex)
PY
}
run default --
run nvls8 SPB_NCCL_NVLS_CTAS=8 --
run nvls16 SPB_NCCL_NVLS_CTAS=16 --
run nvls32 SPB_NCCL_NVLS_CTAS=32 --
run max16 SPB_NCCL_MAX_CTAS=16 --
run tiles1 SPB_GEMM_SHARED_TILES=1 --
run tiles4 SPB_GEMM_SHARED_TILES=4 --
