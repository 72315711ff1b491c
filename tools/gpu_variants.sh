#!/bin/bash
# time tools/diag_flash.py (forward only) with each library variant in solvingpapers_b200/_ab/ named on the command line
mkdir -p gpurun_out
for v in "$@"; do
  lib=$PWD/solvingpapers_b200/_ab/libspb200_$v.so
  [ "$v" = "cur" ] && lib=$PWD/solvingpapers_b200/libspb200.so
  echo "== $v"
  SPB_LIB_PATH=$lib timeout 300 python tools/diag_flash.py > gpurun_out/diag_flash_$v.log 2>&1; grep -E "time fwd|FAIL|EXC|ALL" gpurun_out/diag_flash_$v.log
done
