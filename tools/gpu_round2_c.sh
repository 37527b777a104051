#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python tools/q17_probe.py C3 60000 > $O/r02c_q17_C3_60k.txt 2>&1; cat $O/r02c_q17_C3_60k.txt
python tools/q17_probe.py C2 > $O/r02c_q17_C2.txt 2>&1; cat $O/r02c_q17_C2.txt
python -m pytest tests -m gpu -q --durations=8 > $O/r02c_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02c_pytest_gpu.log
tail -15 $O/r02c_pytest_gpu.log
D=vings-mono_b200/diff_surfel_rasterization
for c in C2 C4 C5; do
  for v in "" _nomask _blockcull; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so python tools/time_stages.py $c 2 >> $O/r02c_time_stages.txt 2>&1
  done
done
cat $O/r02c_time_stages.txt
