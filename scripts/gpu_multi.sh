#!/bin/bash
# N-GPU visit: T3 tests (kdb_reduce_nccl under torchrun, kdb_reduce_peers in one process, CLI --devices) and the N-GPU bench.
# usage: gpurun --gpus N --timeout 1500 -- 'bash scripts/gpu_multi.sh TAG N [reads]'
TAG=$1; N=$2; READS=${3:-0}
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q --timeout 600 > $O/pytest_multi_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_multi_$TAG.log
tail -15 $O/pytest_multi_$TAG.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline ${READS:+--reads $READS} > $O/bench_n${N}_$TAG.json 2> $O/bench_n${N}_$TAG.err; echo "bench exit $?"
tail -5 $O/bench_n${N}_$TAG.err
python - <<PY
import json
d=json.load(open('$O/bench_n${N}_$TAG.json'))
print('N=$N value %.4g ms %.3f wall %.3f' % (d['value'], d['ms_per_step'], d['wall_ms_per_step']), d['kernel_ms_per_step'], 'e2e %.4g' % d['e2e']['value'], d['result_check'])
PY
