#!/bin/bash
# quick look: one bench run (no CPU legs) and the numbers that matter.  usage: gpu_quick.sh TAG [bench args]
TAG=${1:-x}; shift; O=gpurun_out; mkdir -p $O
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" > $O/bench_quick_$TAG.json 2> $O/bench_quick_$TAG.err; echo "bench exit $?"; tail -3 $O/bench_quick_$TAG.err
python - <<PY
import json
d=json.load(open('$O/bench_quick_$TAG.json'))
print('value %.4g ms %.3f e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['kernel_ms_per_step'])
print('bgzf', json.dumps(d.get('e2e_bgzf')))
PY
