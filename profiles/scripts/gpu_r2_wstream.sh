#!/bin/bash
# micro-benchmark of the weight-stage stream (profiles/ubench/wstream.cu) on 1 x B200
set -u
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/wstream profiles/ubench/wstream.cu -lcudart || exit 1
timeout 300 /tmp/wstream > gpurun_out/wstream_r2.txt 2>&1; echo "exit $?"
cat gpurun_out/wstream_r2.txt
