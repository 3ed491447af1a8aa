#!/bin/bash
# Final evidence visit of round 2 (1 GPU): whole parity suite, both bench arms, launch list, inflate probes + ncu of the inflate kernels.
TAG=${1:-r02x}
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q --timeout 1500 > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
tail -4 $O/pytest_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err; echo "bench ref exit $?"
python - <<PY
import json
d=json.load(open('$O/bench_$TAG.json')); r=json.load(open('$O/bench_ref_$TAG.json'))
print('value %.4g ms %.3f e2e %.4g bgzf %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value'], d.get('e2e_bgzf',{}).get('value',0)), d['kernel_ms_per_step'])
print('roofline', d['roofline']['kernel'], d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'], d.get('cpu_baseline_port',{}).get('value'))
print('bgzf', json.dumps(d.get('e2e_bgzf')))
print('ref arm', r['value'], r['cpu_baseline']['kind'], r['cpu_baseline']['cores'])
PY
SHORT="timeout 300 python bench.py --steps 2 --warmup 3 --reads 2000000 --no-cpu-baseline"
$SHORT > $O/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_$TAG.csv $SHORT > $O/ncu_l_$TAG.log 2>&1; echo "launch list exit $?"
for lv in 6 1; do timeout 300 python scripts/inflate_probe.py 10000000 $lv > $O/inflate_${TAG}_l$lv.json 2> $O/inflate_${TAG}_l$lv.err; echo "probe l$lv exit $?"; cat $O/inflate_${TAG}_l$lv.json; done
bash scripts/ncu_inflate.sh $TAG 10000000
