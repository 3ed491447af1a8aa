#!/bin/bash
# ncu --set full of the inflate kernels on the probe (the probe ran clean in gpu_inflate.sh).  usage: ncu_inflate.sh TAG [reads]
TAG=${1:-x}; N=${2:-5000000}; O=gpurun_out; mkdir -p $O
timeout 300 python scripts/inflate_probe.py $N 6 > $O/inflate_plain_$TAG.json 2>&1 || { echo "plain failed"; tail -5 $O/inflate_plain_$TAG.json; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_inflate -c 2 -o $O/prof_inflate_$TAG python scripts/inflate_probe.py $N 6 > $O/ncu_inflate_$TAG.log 2>&1
echo "ncu exit $?"; ls -la $O/prof_inflate_$TAG.ncu-rep
