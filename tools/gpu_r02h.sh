#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
run() {
  name=$1; shift
  envs=(); while [ "$1" != "--" ] && [ $# -gt 0 ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --sustain-s 0 "$@" > gpurun_out/r02h_n${N}_$name.json 2> gpurun_out/r02h_n${N}_$name.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r02h_n${N}_$name.json") if l.startswith("{")][-1])
    s=d["stages_ms"]
    print("  N=$N $name: ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), "| dx", s["gemm_dx"], "allreduce", s["allreduce"], "flash_bwd", s["flash_bwd"], "dwqkv", s["gemm_dwqkv"], "norm_bwd", s["norm_bwd"], "| grads:", d["config"].get("grad_buffers"), "| dp_check", d.get("dp_check",{}).get("max_rel_err"))
except Exception as ex:
    print("  parse failed:", ex)
PY
}
run static SPB_GEMM_SHARED_TILES=0 --
run dyn1 SPB_GEMM_SHARED_TILES=1 --
run dyn2 SPB_GEMM_SHARED_TILES=2 --
run dyn1_noprio SPB_GEMM_SHARED_TILES=1 SPB_COMM_PRIORITY=0 --
tail -3 gpurun_out/r02h_n${N}_dyn1.err
