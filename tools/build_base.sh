#!/bin/bash
# development only: build the library from the sources of a git revision (default HEAD) into solvingpapers_b200/_ab/libspb200_base.so
set -e
rev=${1:-HEAD}
root="$(cd "$(dirname "$0")/.." && pwd)"
rm -rf /tmp/spb_base && mkdir -p /tmp/spb_base "$root/solvingpapers_b200/_ab"
git -C "$root" archive "$rev" solvingpapers_b200/csrc include | tar -x -C /tmp/spb_base
cd /tmp/spb_base/solvingpapers_b200/csrc
for f in api norm_rope gemm_simt gemm_tc attn_simt flash_tc flash_bwd_tc ffn nll comm; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -I../../include -c $f.cu -o /tmp/spb_base/$f.o &
done
wait
nvcc -shared -o "$root/solvingpapers_b200/_ab/libspb200_base.so" /tmp/spb_base/*.o -cudart static -ldl -lpthread -lrt 2>/dev/null
ls -la "$root/solvingpapers_b200/_ab/libspb200_base.so"
