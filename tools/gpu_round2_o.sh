#!/bin/bash
# Ring gather by cp.async (LDGSTS, default build) against the TMA bulk gather (variant tma): timing first on the config
# that is bound by it, then the full suite on the new default, then the rest as far as the GPU budget reaches.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
D=$PWD/vings-mono_b200/diff_surfel_rasterization
run() { SB200_LIB=$D/libsurfel_b200$2.so timeout 100 python tools/time_stages.py $1 2 >> $O/r02o_time_stages.txt 2>&1; }
run C5 ""; run C5 _tma
cut -c1-30,95-230 $O/r02o_time_stages.txt
timeout 280 python -m pytest tests -m gpu -x -q > $O/r02o_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02o_pytest_gpu.log; tail -3 $O/r02o_pytest_gpu.log
run C2 ""; run C2 _tma
cut -c1-30,95-230 $O/r02o_time_stages.txt
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02o_smoke.log 2>&1; echo "smoke_rc=$?" >> $O/r02o_smoke.log; tail -2 $O/r02o_smoke.log
timeout 150 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o $O/r02o_view_C2 \
    python tools/profile_view.py C2 > $O/r02o_ncu_view.log 2>&1; tail -1 $O/r02o_ncu_view.log
cp $O/view_C2.csrc_hash $O/r02o_view_C2.csrc_hash 2>/dev/null
run C4 ""; run C4 _tma
timeout 200 python bench.py > $O/r02o_bench_b200_C2.json 2> $O/r02o_bench_b200_C2.err; echo "bench_rc=$?"
