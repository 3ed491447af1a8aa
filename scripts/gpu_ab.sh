#!/bin/bash
# A/B visit: GPU parity tests, then short benches of the variants named by env assignments (one per argument).
# usage: gpurun --timeout 1200 -- 'bash scripts/gpu_ab.sh TAG "" "KDB_TWO_SWEEP=1"'
TAG=$1; shift
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 ${PYTEST_K:+-k "$PYTEST_K"} > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
tail -4 $O/pytest_$TAG.log
i=0
for V in "$@"; do
  env $V timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_${TAG}_$i.json 2> $O/bench_${TAG}_$i.err; echo "bench[$V] exit $?"
  python - <<PY
import json
d=json.load(open('$O/bench_${TAG}_$i.json'))
print('$V', 'value %.4g' % d['value'], 'ms %.3f' % d['ms_per_step'], d['kernel_ms_per_step'], 'e2e %.4g' % d['e2e']['value'])
PY
  i=$((i+1))
done
