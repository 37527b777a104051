#!/bin/bash
# A/B: per-pixel constants of the backward blend in shared memory (BB_PIX_N = 5 static, 10 dynamic, 10 with 4 CTAs/SM everywhere)
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
D=vings-mono_b200/diff_surfel_rasterization
SB200_LIB=$PWD/$D/libsurfel_b200_pix10.so timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_ops_gpu.py -m gpu -x -q > $O/r02j_pytest_pix10.log 2>&1; echo "pytest_rc=$?" >> $O/r02j_pytest_pix10.log; tail -3 $O/r02j_pytest_pix10.log
for c in C2 C4 C5; do
  for v in "" _pix5 _pix10 _pix10c4; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so timeout 300 python tools/time_stages.py $c 2 >> $O/r02j_time_stages.txt 2>&1
  done
done
cat $O/r02j_time_stages.txt
