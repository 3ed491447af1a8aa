#!/bin/bash
# ncu --set full of the kernels matching REGEX in one short bench run (after the same command ran clean without ncu)
# usage: bash scripts/ncu_any.sh TAG REGEX [skip] [count] [bench args...]
TAG=${1:-x}; RE=${2:-k2s_}; SKIP=${3:-3}; CNT=${4:-2}; shift 4
O=gpurun_out; mkdir -p $O
SHORT="timeout 300 python bench.py --steps 2 --warmup 3 --reads 2000000 --no-cpu-baseline $@"
$SHORT > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TAG.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c $CNT -o $O/prof_$TAG $SHORT > $O/ncu_f_$TAG.log 2>&1
echo "ncu exit $?"; ls -la $O/prof_$TAG.ncu-rep
