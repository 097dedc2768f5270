#!/bin/bash
# micro-benchmark of the weight-ring handshake (profiles/ubench/handshake.cu) on 1 x B200
set -u
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf-replication_b200/csrc -o /tmp/hs profiles/ubench/handshake.cu -lcudart 2>/dev/null || exit 1
timeout 120 /tmp/hs > gpurun_out/handshake_r2.txt 2>&1; echo "exit $?"
cat gpurun_out/handshake_r2.txt
