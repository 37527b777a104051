#!/bin/bash
# A/B: dLc in shared memory too (n3_13), a 6-stage ring (st6), 6 stages + all 10 constants at 4 CTAs/SM (st6_n4_10)
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
D=vings-mono_b200/diff_surfel_rasterization
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_ops_gpu.py -m gpu -x -q > $O/r02l_pytest.log 2>&1; echo "pytest_rc=$?" >> $O/r02l_pytest.log; tail -3 $O/r02l_pytest.log
for c in C2 C4 C5; do
  for v in "" _n3_13 _st6 _st6_n4_10; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so timeout 300 python tools/time_stages.py $c 2 >> $O/r02l_time_stages.txt 2>&1
  done
done
cut -c1-40,120-200 $O/r02l_time_stages.txt
