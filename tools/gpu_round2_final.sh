#!/bin/bash
# Final 1-GPU evidence run of round 2: the suite three times, smoke, both bench arms, ncu captures.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
for i in 1 2 3; do
  python -m pytest tests -m gpu -x -q -s > $O/r02_final_pytest_gpu_$i.log 2>&1; echo "pytest_rc=$?" >> $O/r02_final_pytest_gpu_$i.log; tail -2 $O/r02_final_pytest_gpu_$i.log
done
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_final_smoke.log 2>&1; echo "smoke_rc=$?" >> $O/r02_final_smoke.log; tail -2 $O/r02_final_smoke.log
python bench.py > $O/r02_final_bench_b200_C2.json 2> $O/r02_final_bench_b200_C2.err
python bench.py --impl reference > $O/r02_final_bench_reference_C2.json 2> $O/r02_final_bench_reference_C2.err
python bench.py --config C3 --no-extras > $O/r02_final_bench_b200_C3.json 2> $O/r02_final_bench_b200_C3.err
python bench.py --config C5 --steps 3 --warmup 3 --no-extras > $O/r02_final_bench_b200_C5_churn.json 2> $O/r02_final_bench_b200_C5_churn.err
python bench.py --config C1 --steps 20 --warmup 5 --no-extras > $O/r02_final_bench_b200_C1.json 2> $O/r02_final_bench_b200_C1.err
timeout 900 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o $O/r02_final_view_C2 \
    python tools/profile_view.py C2 > $O/r02_final_ncu_view.log 2>&1; tail -1 $O/r02_final_ncu_view.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02_final_launches_bench_C2.csv \
    python bench.py --steps 2 --warmup 1 --no-extras > $O/r02_final_ncu_bench.log 2>&1
for c in C4 C5; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:sb_blend -c 2 -o $O/r02_final_${c}_blend -f \
      python tools/profile_raster.py $c 1 > $O/r02_final_ncu_$c.log 2>&1
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02_final_bench_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'raster', round(d.get('raster',{}).get('value',0),1),
              'dropin', d.get('dropin',{}).get('value'), 'strong', d.get('strong',{}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
