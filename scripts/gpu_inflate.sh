#!/bin/bash
# inflate visit: tests + probe of both inflate paths.  usage: gpu_inflate.sh TAG [reads]
TAG=${1:-x}; N=${2:-10000000}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_inflate.py -x -q > gpurun_out/pytest_inflate_$TAG.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_inflate_$TAG.log
for lv in 6 1; do
  timeout 300 python scripts/inflate_probe.py $N $lv > gpurun_out/inflate_${TAG}_l$lv.json 2> gpurun_out/inflate_${TAG}_l$lv.err; echo "probe l$lv exit $?"; cat gpurun_out/inflate_${TAG}_l$lv.json; tail -3 gpurun_out/inflate_${TAG}_l$lv.err
done
KDB_INFLATE_WARP=1 timeout 300 python scripts/inflate_probe.py $N 6 > gpurun_out/inflate_${TAG}_warp_l6.json 2>&1; echo "probe warp exit $?"; cat gpurun_out/inflate_${TAG}_warp_l6.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv -k regex:k_inflate --log-file gpurun_out/inflate_launches_$TAG.csv python scripts/inflate_probe.py $N 6 > gpurun_out/inflate_ncu_$TAG.log 2>&1; echo "ncu exit $?"
grep k_inflate gpurun_out/inflate_launches_$TAG.csv | awk -F'","' '{print $5, $NF}' | head -20
