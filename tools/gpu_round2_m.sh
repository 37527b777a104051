#!/bin/bash
# Forward blend with the warp-level interval cull (sb_block_far) and the padded ring pitch: full suite + smoke on the
# new default, then A/B against the previous kernel (old), the cull without the padding (far8) and the backward's
# sparse ring fill (sparse), then the parity files on the sparse variant.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
D=$PWD/vings-mono_b200/diff_surfel_rasterization
timeout 420 python -m pytest tests -m gpu -x -q > $O/r02m_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02m_pytest_gpu.log; tail -3 $O/r02m_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/r02m_smoke.log 2>&1; echo "smoke_rc=$?" >> $O/r02m_smoke.log; tail -2 $O/r02m_smoke.log
run() { SB200_LIB=$D/libsurfel_b200$2.so timeout 120 python tools/time_stages.py $1 2 >> $O/r02m_time_stages.txt 2>&1; }
for c in C2 C5 C4; do for v in "" _old _sparse; do run $c "$v"; done; done
run C2 _far8; run C5 _far8; run C3 ""; run C3 _old
cut -c1-30,95-230 $O/r02m_time_stages.txt
SB200_LIB=$D/libsurfel_b200_sparse.so timeout 300 python -m pytest tests/test_parity_gpu.py tests/test_parity_full_gpu.py tests/test_ops_gpu.py -m gpu -x -q > $O/r02m_pytest_sparse.log 2>&1; echo "pytest_rc=$?" >> $O/r02m_pytest_sparse.log; tail -3 $O/r02m_pytest_sparse.log
timeout 200 python bench.py --no-extras --steps 10 --warmup 3 > $O/r02m_bench_b200_C2_noextras.json 2> $O/r02m_bench.err; tail -c 600 $O/r02m_bench_b200_C2_noextras.json
