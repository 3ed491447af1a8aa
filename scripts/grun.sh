#!/bin/bash
# gpurun with retries while the pod answers "transient" (nothing charged).
# usage: scripts/grun.sh TIMEOUT 'command'            (1 GPU)
#        GPUS=2 scripts/grun.sh TIMEOUT 'command'     (N GPUs of one box)
T=$1; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun ${GPUS:+--gpus $GPUS} --timeout $T -- "$@" > /tmp/grun_last.log 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/grun_last.log || [ $rc -eq 3 ]; then sleep 45; continue; fi
  break
done
cat /tmp/grun_last.log
