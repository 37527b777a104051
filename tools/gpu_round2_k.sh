#!/bin/bash
# The backward blend with its per-pixel constants in shared memory (10 at 3 CTAs/SM, 5 at 4): suite, then A/B of 3 / 5 / 7 at 4 CTAs/SM
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
D=vings-mono_b200/diff_surfel_rasterization
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02k_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02k_pytest_gpu.log; tail -3 $O/r02k_pytest_gpu.log
SB200_LIB=$PWD/$D/libsurfel_b200.so timeout 300 python tools/time_stages.py C2 2 >> $O/r02k_time_stages.txt 2>&1
for c in C4 C5; do
  for v in "" _n4_3 _n4_7; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so timeout 300 python tools/time_stages.py $c 2 >> $O/r02k_time_stages.txt 2>&1
  done
done
cat $O/r02k_time_stages.txt
