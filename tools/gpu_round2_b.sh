#!/bin/bash
# GPU call B of round 2: parity with the forward hit masks, A/B of the forward cull variants, reference-arm lines of
# the other configs, and the in-situ probe.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q --durations=8 > $O/r02b_pytest_gpu.log 2>&1; echo "pytest_rc=$?" >> $O/r02b_pytest_gpu.log
tail -5 $O/r02b_pytest_gpu.log
D=vings-mono_b200/diff_surfel_rasterization
for c in C2 C3 C4 C5; do
  for v in "" _nolive _nocorner _plain; do
    SB200_LIB=$PWD/$D/libsurfel_b200$v.so python tools/time_stages.py $c 2 >> $O/r02b_time_stages.txt 2>&1
  done
done
cat $O/r02b_time_stages.txt
python tools/insitu_probe.py C2 8 > $O/r02b_insitu_C2.txt 2>&1; cat $O/r02b_insitu_C2.txt
python bench.py > $O/r02b_bench_b200_C2.json 2> $O/r02b_bench_b200_C2.err
for c in C3 C4 C5; do
  python bench.py --impl reference --config $c --steps 2 --warmup 3 > $O/r02b_bench_reference_$c.json 2> $O/r02b_bench_reference_$c.err
  python bench.py --config $c --steps 3 --warmup 3 > $O/r02b_bench_b200_$c.json 2> $O/r02b_bench_b200_$c.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02b_bench_*.json')):
    try:
        d = json.load(open(f)); print(f, round(d['value'],1), round(d['e2e']['value'],1), d.get('raster',{}).get('value'))
    except Exception as e: print(f, 'ERR', e)
PY
