#!/bin/bash
# one bench line per BASELINE config (other than the default C3) -> gpurun_out/workloads/
mkdir -p gpurun_out/workloads
for w in c1 c2 c2lit c4 c4lit c5; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/workloads/bench_$w.log 2> gpurun_out/workloads/bench_$w.err; echo "$w rc=$?"
  python tools/show_bench.py gpurun_out/workloads/bench_$w.log
done
