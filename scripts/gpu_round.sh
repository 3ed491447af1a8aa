#!/bin/bash
# One GPU-box visit: parity tests, bench (both arms), ncu launch list and full captures of the top kernels.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh TAG'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $O/clocks_$TAG.csv &
SMI=$!
python -m pytest tests -m gpu -x -q > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log
tail -3 $O/pytest_$TAG.log
python bench.py --steps 20 --warmup 3 > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench exit $?"
cat $O/bench_$TAG.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err
cat $O/bench_ref_$TAG.json
kill $SMI
SHORT="python bench.py --steps 2 --warmup 3 --reads 2000000 --no-cpu-baseline"
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$TAG.csv $SHORT > $O/ncu_l_$TAG.log 2>&1
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k3p_|k2_encode|k1_tile' -s 9 -c 4 -o $O/prof_$TAG $SHORT > $O/ncu_f_$TAG.log 2>&1
ls -la $O
