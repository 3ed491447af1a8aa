#!/bin/bash
# N-GPU evidence: T3 tests, bench at N (configs[2]), and configs[4] (graph + minimizers over all 100 M reads).
TAG=$1; N=$2
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q --timeout 600 > $O/pytest_multi_n${N}_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_multi_n${N}_$TAG.log
tail -3 $O/pytest_multi_n${N}_$TAG.log
for M in $3; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $M --master-addr 127.0.0.1 --master-port 2957$M bench.py --gpus $M --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_n${M}_$TAG.json 2> $O/bench_n${M}_$TAG.err; echo "bench N=$M exit $?"
  python - <<PY
import json
d=json.load(open('$O/bench_n${M}_$TAG.json'))
print('N=$M value %.4g ms %.3f wall %.3f' % (d['value'], d['ms_per_step'], d['wall_ms_per_step']), d['kernel_ms_per_step'], 'e2e %.4g' % d['e2e']['value'], d['result_check'].get('t3_identical'))
PY
done
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29588 scripts/config5_graph_min.py > $O/config5_n${N}_$TAG.log 2>&1; echo "config5 exit $?"
tail -4 $O/config5_n${N}_$TAG.log | cut -c1-900
