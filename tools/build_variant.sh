#!/bin/bash
# development only: build the library with extra nvcc flags into solvingpapers_b200/_ab/libspb200_<name>.so
#   tools/build_variant.sh trace -DSPB_FA_TRACE        tools/build_variant.sh poly1 -DSPB_FA_POLY_EVERY=1
set -e
name=$1; shift
cd "$(dirname "$0")/../solvingpapers_b200/csrc"
mkdir -p ../_ab /tmp/spb_$name
for f in api norm_rope gemm_simt gemm_tc attn_simt flash_tc flash_bwd_tc ffn nll comm decode attn_decode_mma decode_gemm_mma attn_wide attn_ffma gemm_split; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -I../../include -c $f.cu -o /tmp/spb_$name/$f.o &
done
wait
nvcc -shared -o ../_ab/libspb200_$name.so /tmp/spb_$name/*.o -cudart static -ldl -lpthread -lrt 2>/dev/null
ls -la ../_ab/libspb200_$name.so
