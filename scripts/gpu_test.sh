#!/bin/bash
# quick GPU-box visit: parity tests (optionally a -k filter) and a short bench
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q ${1:+-k "$1"} > $O/pytest_quick.log 2>&1; echo "pytest exit $?" >> $O/pytest_quick.log
tail -25 $O/pytest_quick.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_quick.json 2> $O/bench_quick.err; echo "bench exit $?"; tail -3 $O/bench_quick.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print(d['value'], d['ms_per_step'], d['kernel_ms_per_step'], d['e2e']['value'])
PY
