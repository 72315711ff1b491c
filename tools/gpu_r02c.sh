#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py -q --timeout 600 -p no:cacheprovider -x > gpurun_out/r02e_tests_decode.log 2>&1; echo "decode pytest rc=$?"
tail -n 8 gpurun_out/r02e_tests_decode.log
for wl in c3dec c3dec32; do
  timeout 600 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r02e_bench_$wl.json 2> gpurun_out/r02e_bench_$wl.err; echo "bench $wl rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r02e_bench_$wl.json"))
r=d["roofline"]
print("$wl", "ms/step", round(d["ms_per_step"],4), "step GB/s", round(r["achieved"]), "frac", round(r["frac"],3), "attn ms", round(r["attn_decode_kernel"]["ms"],4), "attn GB/s", round(r["attn_decode_kernel"]["gbs"]), "e2e ms", round(d["e2e"]["ms_per_step"],4))
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02e_launches_c3dec.csv python bench.py --workload c3dec --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02e_ncu.log 2>&1; echo "ncu rc=$?"
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/r02e_launches_c3dec.csv")) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if "Kernel Name" in r][0]
h=rows[hdr]; ki=h.index("Kernel Name"); vi=h.index("Metric Value")
agg=collections.defaultdict(list)
for r in rows[hdr+1:]:
    try: agg[r[ki].split("(")[0][:60]].append(float(r[vi].replace(",","")))
    except Exception: pass
for k,v in sorted(agg.items(), key=lambda kv:-sum(kv[1])):
    print(f"{k:62s} n={len(v):3d} avg {sum(v)/len(v)/1e3:8.2f} us")
PY
