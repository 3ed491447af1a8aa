#!/bin/bash
# the no-alloc entry point: its test alone in a fresh process, then the whole GPU suite (which collects it last)
TAG=${1:-x}; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_noalloc_alone_$TAG.log 2>&1; echo "alone exit $?"; tail -4 gpurun_out/pytest_noalloc_alone_$TAG.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "suite exit $?"; tail -4 gpurun_out/pytest_gpu_$TAG.log
