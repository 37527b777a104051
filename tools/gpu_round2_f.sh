#!/bin/bash
# N-GPU call: multi-rank parity + bench through both exchanges
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
N=${1:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29521 tools/dp_check.py > $O/r02f_dp_check_${N}gpu.json 2> $O/r02f_dp_check_${N}gpu.err; echo "dp_check rc=$?"; tail -1 $O/r02f_dp_check_${N}gpu.json
timeout 600 $RUN --master-port 29522 bench.py --gpus $N --steps 5 --warmup 3 > $O/r02f_bench_b200_C2_${N}gpu.json 2> $O/r02f_bench_b200_C2_${N}gpu.err; echo "bench p2p rc=$?"; tail -3 $O/r02f_bench_b200_C2_${N}gpu.err
[ -n "$SKIP_NCCL" ] || SB200_DP_EXCHANGE=nccl timeout 600 $RUN --master-port 29523 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $O/r02f_bench_b200_C2_${N}gpu_nccl.json 2> $O/r02f_bench_b200_C2_${N}gpu_nccl.err; echo "bench nccl rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02f_bench_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'identical', d.get('replicas_bit_identical'),
              'strong', d.get('strong',{}).get('value'), {k: round(v['ms_per_launch'],3) for k,v in d['stages'].items() if k in ('update','dp_reduce','dp_post','preprocess_bwd')})
    except Exception as e: print(f, 'ERR', e)
PY
