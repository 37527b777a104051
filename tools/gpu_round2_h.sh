#!/bin/bash
# 8-GPU diagnostics of the exchange: wait accounting, no-skew schedule, no host confirm
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
N=${1:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $RUN --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $O/r02h_default_${N}gpu.json 2> $O/r02h_default_${N}gpu.err
SB200_BENCH_SAME_VIEW=1 timeout 400 $RUN --master-port 29532 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $O/r02h_sameview_${N}gpu.json 2> $O/r02h_sameview_${N}gpu.err
SB200_DIAG_NO_CONFIRM=1 timeout 400 $RUN --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $O/r02h_noconfirm_${N}gpu.json 2> $O/r02h_noconfirm_${N}gpu.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02h_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d.get('dp_wait_ms_per_view'), {k: round(v['ms_per_launch'],3) for k,v in d['stages'].items() if k in ('update','dp_reduce','dp_post','blend_bwd','blend_fwd')})
    except Exception as e: print(f, 'ERR', e)
PY
