#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q --durations=6 > $O/r02d_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02d_pytest_gpu.log
tail -12 $O/r02d_pytest_gpu.log
D=vings-mono_b200/diff_surfel_rasterization
for c in C2 C4 C5; do
  for v in "" _nodiag _cta2 _cta4; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so python tools/time_stages.py $c 2 >> $O/r02d_time_stages.txt 2>&1
  done
done
cat $O/r02d_time_stages.txt
python bench.py > $O/r02d_bench_b200_C2.json 2> $O/r02d_bench_b200_C2.err; tail -3 $O/r02d_bench_b200_C2.err
python bench.py --impl reference --steps 3 --warmup 3 > $O/r02d_bench_reference_C2.json 2> $O/r02d_bench_reference_C2.err; tail -3 $O/r02d_bench_reference_C2.err
python bench.py --config C5 --steps 3 --warmup 3 --no-extras > $O/r02d_bench_b200_C5_churn.json 2> $O/r02d_bench_b200_C5_churn.err; tail -3 $O/r02d_bench_b200_C5_churn.err
python bench.py --config C5 --impl reference --steps 2 --warmup 3 --no-extras > $O/r02d_bench_reference_C5_churn.json 2> $O/r02d_bench_reference_C5_churn.err; tail -3 $O/r02d_bench_reference_C5_churn.err
python bench.py --config C1 --steps 20 --warmup 5 --no-extras > $O/r02d_bench_b200_C1.json 2> $O/r02d_bench_b200_C1.err
python bench.py --config C1 --impl reference --steps 20 --warmup 5 --no-extras > $O/r02d_bench_reference_C1.json 2> $O/r02d_bench_reference_C1.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02d_bench_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'raster', round(d.get('raster',{}).get('value',0),1),
              'dropin', d.get('dropin',{}).get('value'), 'strong', d.get('strong',{}).get('value'), 'churn', d.get('churn',{}).get('surfels_per_timed_step'))
    except Exception as e: print(f, 'ERR', e)
PY
