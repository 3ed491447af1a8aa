#!/bin/bash
# ncu --set full of the partition + count kernels of one short bench run (after the same command ran clean without ncu)
TAG=${1:-x}
O=gpurun_out; mkdir -p $O
SHORT="python bench.py --steps 2 --warmup 3 --reads 2000000 --no-cpu-baseline"
$SHORT > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TAG.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'k3p_|k3s_' -s 6 -c 2 -o $O/prof_$TAG $SHORT > $O/ncu_f_$TAG.log 2>&1
echo "ncu exit $?"; ls -la $O/prof_$TAG.ncu-rep
