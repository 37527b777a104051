#!/bin/bash
# GPU call A of round 2: the parity suite (looped for flakiness), smoke, a first bench line, stage times of the
# large configs and ncu captures of one C4 / C5 view (baseline before the kernel work).
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q --durations=12 > $O/r02_pytest_gpu_1.log 2>&1; echo "pytest_rc=$?" >> $O/r02_pytest_gpu_1.log
python -m pytest tests -m gpu -q > $O/r02_pytest_gpu_2.log 2>&1; echo "pytest_rc=$?" >> $O/r02_pytest_gpu_2.log
for i in 3 4 5; do
  python -m pytest tests -m gpu -q --ignore=tests/test_parity_full_gpu.py > $O/r02_pytest_gpu_$i.log 2>&1; echo "pytest_rc=$?" >> $O/r02_pytest_gpu_$i.log
done
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke.log 2>&1; echo "smoke_rc=$?" >> $O/r02_smoke.log
python bench.py > $O/r02a_bench_b200_C2.json 2> $O/r02a_bench_b200_C2.err
for c in C2 C4 C5; do python tools/time_stages.py $c 2 >> $O/r02a_time_stages.txt 2>&1; done
for c in C4 C5; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:sb_blend -c 2 -o $O/r02a_${c}_blend -f \
      python tools/profile_raster.py $c 1 > $O/r02a_ncu_$c.log 2>&1
done
tail -3 $O/r02_pytest_gpu_*.log
cat $O/r02a_time_stages.txt
