#!/bin/bash
# ncu --set full of the count-path kernels at one k (after the same command ran clean without ncu)
K=${1:-14}; TAG=${2:-x}
O=gpurun_out; mkdir -p $O
CMD="python scripts/sweep_k.py 2000000 $K $K"
$CMD > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TAG.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'k3p_' -s 4 -c 2 -o $O/prof_$TAG $CMD > $O/ncu_f_$TAG.log 2>&1
echo "ncu exit $?"; ls -la $O/prof_$TAG.ncu-rep
