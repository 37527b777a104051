#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q --durations=6 -s > $O/r02e_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02e_pytest_gpu.log
grep "excluded set\|Q17 excludes" $O/r02e_pytest_gpu.log; tail -6 $O/r02e_pytest_gpu.log
python bench.py > $O/r02e_bench_b200_C2.json 2> $O/r02e_bench_b200_C2.err; tail -3 $O/r02e_bench_b200_C2.err
python tools/time_stages.py C2 2 > $O/r02e_time_stages.txt 2>&1; cat $O/r02e_time_stages.txt
# ncu: one mapping view, every kernel (traffic.json); then the launch list of the bench command
timeout 900 ncu --set full --import-source on --clock-control none --profile-from-start off -f -o $O/r02e_view_C2 \
    python tools/profile_view.py C2 > $O/r02e_ncu_view.log 2>&1; tail -2 $O/r02e_ncu_view.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r02e_launches_bench_C2.csv \
    python bench.py --steps 2 --warmup 1 --no-extras > $O/r02e_ncu_bench.log 2>&1; tail -1 $O/r02e_ncu_bench.log | cut -c1-200
python - <<'PY'
import json
d = json.load(open('gpurun_out/r02e_bench_b200_C2.json'))
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'raster', round(d['raster']['value'],1), 'dropin', d.get('dropin',{}).get('value'), 'strong', d.get('strong',{}).get('value'))
print({k: round(v['ms_per_launch'],4) for k,v in d['stages'].items()}, d['gpu_launches'])
print(d['roofline'])
PY
