#!/bin/bash
# last visit of the round: smoke(), the whole GPU suite, the default bench line (both arms) with wall times
TAG=${1:-last}; O=gpurun_out; mkdir -p $O
t0=$(date +%s); timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1; echo "smoke exit $? ($(( $(date +%s) - t0 )) s)"; tail -3 $O/smoke_$TAG.log
t0=$(date +%s); timeout 2400 python -m pytest tests -x -q -m gpu > $O/pytest_$TAG.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - t0 )) s)"; tail -3 $O/pytest_$TAG.log
t0=$(date +%s); timeout 900 python bench.py > $O/bench_default_$TAG.json 2> $O/bench_default_$TAG.err; echo "bench exit $? ($(( $(date +%s) - t0 )) s)"
t0=$(date +%s); timeout 900 python bench.py --impl reference > $O/bench_default_ref_$TAG.json 2>> $O/bench_default_$TAG.err; echo "bench ref exit $? ($(( $(date +%s) - t0 )) s)"
python - <<PY
import json
d=json.load(open('$O/bench_default_$TAG.json')); r=json.load(open('$O/bench_default_ref_$TAG.json'))
print('value %.4g ms %.3f e2e %.4g bgzf %.4g steps %s warmup %s launches %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d.get('e2e_bgzf',{}).get('value',0), d['steps'], d['warmup'], d['gpu_launches']))
print('ref', r['value'], r.get('impl'))
PY
