#!/bin/bash
# multi-GPU experiments: bash tools/gpu_r02g.sh N
N=${1:-2}
mkdir -p gpurun_out
run() {  # name, env..., -- extra bench args
  name=$1; shift
  envs=(); while [ "$1" != "--" ] && [ $# -gt 0 ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --sustain-s 0 "$@" > gpurun_out/r02g_n${N}_$name.json 2> gpurun_out/r02g_n${N}_$name.err
  echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/r02g_n${N}_$name.json") if l.startswith("{")][-1])
    s=d["stages_ms"]
    print("  N=$N $name: ms/step", round(d["ms_per_step"],4), "Mtok/s", round(d["value"]/1e6,3), "| dx", s["gemm_dx"], "allreduce", s["allreduce"], "flash_bwd", s["flash_bwd"], "dwqkv", s["gemm_dwqkv"], "norm_bwd", s["norm_bwd"], "| grads:", d["config"].get("grad_buffers"), "| dp_check", d.get("dp_check",{}).get("max_rel_err"))
except Exception as ex:
    print("  parse failed:", ex)
PY
}
run base -- --reg-grads off
run reg -- --reg-grads on
run reg_eff SPB_NCCL_CTA_POLICY=1 -- --reg-grads on
run reg_nowin SPB_NCCL_WINDOW=0 -- --reg-grads on
run base_eff SPB_NCCL_CTA_POLICY=1 -- --reg-grads off
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL,TUNING,REG timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e --sustain-s 0 --reg-grads on > gpurun_out/r02g_n${N}_nccl_debug.log 2>&1
grep -E "NVLS|Connected|AllReduce.*(algo|Algo)|Channel|window|Register|register" gpurun_out/r02g_n${N}_nccl_debug.log | grep -v "Ring [0-9]" | sort | uniq -c | sort -rn | head -40 > gpurun_out/r02g_n${N}_nccl_summary.txt
head -30 gpurun_out/r02g_n${N}_nccl_summary.txt
tail -3 gpurun_out/r02g_n${N}_reg.err
