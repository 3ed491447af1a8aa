#!/bin/bash
# Round-2 evidence visit (1 GPU): full parity suite, both bench arms, launch list + full ncu capture, probes.
TAG=${1:-r02}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $O/clocks_$TAG.csv &
SMI=$!
timeout 2400 python -m pytest tests -m gpu -q --timeout 1500 > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
tail -4 $O/pytest_$TAG.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err; echo "bench ref exit $?"
kill $SMI
python - <<PY
import json
d=json.load(open('$O/bench_$TAG.json')); r=json.load(open('$O/bench_ref_$TAG.json'))
print('value %.4g ms %.3f e2e %.4g bgzf %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value'], d.get('e2e_bgzf',{}).get('value',0)), d['kernel_ms_per_step'])
print('roofline', d['roofline']['kernel'], d['roofline']['frac'], 'cpu', d['cpu_baseline']['value'], d.get('cpu_baseline_port',{}).get('value'))
print('ref arm', r['value'], r['cpu_baseline']['kind'], r['cpu_baseline']['cores'])
PY
SHORT="timeout 300 python bench.py --steps 2 --warmup 3 --reads 2000000 --no-cpu-baseline --no-bgzf"
$SHORT > $O/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv $SHORT > $O/ncu_l_$TAG.log 2>&1
$SHORT > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k2s_|k3s_|k3p_|k4_stitch|k3_fold' -s 12 -c 5 -o $O/prof_$TAG $SHORT > $O/ncu_f_$TAG.log 2>&1
for s in sweep_k fastq_probe skew_probe; do timeout 600 python scripts/$s.py > $O/${s}_$TAG.log 2>&1; echo "$s exit $?"; tail -2 $O/${s}_$TAG.log | cut -c1-400; done
ls $O | head -50
