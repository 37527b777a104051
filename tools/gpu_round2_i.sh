#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r02i_pytest_gpu.log 2>&1; tail -3 $O/r02i_pytest_gpu.log
python bench.py --no-extras > $O/r02i_bench_b200_C2.json 2> $O/r02i_bench_b200_C2.err
python - <<'PY'
import json
d = json.load(open('gpurun_out/r02i_bench_b200_C2.json'))
print('value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'raster', round(d['raster']['value'],1))
print({k: round(v['ms_per_launch'],4) for k,v in d['stages'].items()}, d['gpu_launches'])
PY
