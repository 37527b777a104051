#!/bin/bash
# 2-GPU call: multi-rank parity (tools/dp_check.py), the torchrun test, bench at 2 GPUs through both exchanges.
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
mkdir -p $O
N=${1:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/r02e_topo_${N}gpu.txt 2>&1
timeout 300 $RUN --master-port 29511 tools/dp_check.py > $O/r02e_dp_check_${N}gpu.json 2> $O/r02e_dp_check_${N}gpu.err; echo "dp_check rc=$?"; cat $O/r02e_dp_check_${N}gpu.json; tail -5 $O/r02e_dp_check_${N}gpu.err
python -m pytest tests/test_mapping.py tests/test_multigpu.py -m gpu -q -x > $O/r02e_pytest_multigpu.log 2>&1; tail -3 $O/r02e_pytest_multigpu.log
timeout 600 $RUN --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 > $O/r02e_bench_b200_C2_${N}gpu.json 2> $O/r02e_bench_b200_C2_${N}gpu.err; echo "bench p2p rc=$?"; tail -3 $O/r02e_bench_b200_C2_${N}gpu.err
SB200_DP_EXCHANGE=nccl timeout 600 $RUN --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --no-extras > $O/r02e_bench_b200_C2_${N}gpu_nccl.json 2> $O/r02e_bench_b200_C2_${N}gpu_nccl.err; echo "bench nccl rc=$?"; tail -3 $O/r02e_bench_b200_C2_${N}gpu_nccl.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r02e_bench_*.json')):
    try:
        d = json.load(open(f))
        print(f, 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'identical', d.get('replicas_bit_identical'),
              'strong', d.get('strong',{}).get('value'), d['config']['parallelism'][:90])
    except Exception as e: print(f, 'ERR', e)
PY
