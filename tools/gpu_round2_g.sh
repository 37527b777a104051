#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python tools/time_stages.py C2 2 --per-view > $O/r02g_per_view_C2.txt 2>&1; cat $O/r02g_per_view_C2.txt
D=vings-mono_b200/diff_surfel_rasterization
for c in C2 C3 C4 C5; do
  for v in "" _poll200 _poll1000; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so python tools/time_stages.py $c 2 >> $O/r02g_time_stages.txt 2>&1
  done
done
cat $O/r02g_time_stages.txt
python -m pytest tests -m gpu -x -q > $O/r02g_pytest_gpu.log 2>&1; tail -3 $O/r02g_pytest_gpu.log
