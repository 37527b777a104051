#!/bin/bash
# Final 1-GPU evidence of round 2 after the forward interval cull / sparse backward fill: suite, smoke, the default bench
# line, ncu captures of the final kernels (one C2 mapping view for traffic.json; the two blend kernels on a C5 view),
# the C5 churn and C3 lines, the launch list.  Ordered by priority: the call may be cut short by the GPU budget.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > $O/r02n_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02n_pytest_gpu.log; tail -3 $O/r02n_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02n_smoke.log 2>&1; echo "smoke_rc=$?" >> $O/r02n_smoke.log; tail -2 $O/r02n_smoke.log
timeout 300 python bench.py > $O/r02n_bench_b200_C2.json 2> $O/r02n_bench_b200_C2.err; echo "bench_rc=$?"
timeout 200 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o $O/r02n_view_C2 \
    python tools/profile_view.py C2 > $O/r02n_ncu_view.log 2>&1; tail -1 $O/r02n_ncu_view.log
cp $O/view_C2.csrc_hash $O/r02n_view_C2.csrc_hash 2>/dev/null
python tools/ncu_summary.py $O/r02n_view_C2.ncu-rep > $O/r02n_view_C2_ncu_summary.txt 2>&1
timeout 150 ncu --set full --import-source on --clock-control none -k regex:sb_blend -c 2 -o $O/r02n_C5_blend -f \
    python tools/profile_raster.py C5 1 > $O/r02n_ncu_C5.log 2>&1
python tools/ncu_summary.py $O/r02n_C5_blend.ncu-rep > $O/r02n_C5_blend_ncu_summary.txt 2>&1; cat $O/r02n_C5_blend_ncu_summary.txt
timeout 150 python bench.py --config C5 --steps 3 --warmup 3 --no-extras > $O/r02n_bench_b200_C5_churn.json 2> $O/r02n_bench_b200_C5_churn.err
timeout 100 python bench.py --config C3 --no-extras > $O/r02n_bench_b200_C3.json 2> $O/r02n_bench_b200_C3.err
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02n_launches_bench_C2.csv \
    python bench.py --steps 2 --warmup 1 --no-extras > $O/r02n_ncu_bench.log 2>&1
timeout 100 python bench.py --config C1 --steps 20 --warmup 5 --no-extras > $O/r02n_bench_b200_C1.json 2> $O/r02n_bench_b200_C1.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02n_bench_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'raster', round(d.get('raster',{}).get('value',0),1),
              'dropin', d.get('dropin',{}).get('value'), 'strong', d.get('strong',{}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
ls -la $O/*.ncu-rep
