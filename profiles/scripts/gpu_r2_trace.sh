#!/bin/bash
# event traces of the saving forward's CTA 0: product flavour and the "no MMAs, 1 KB weight stages" flavour (pure handshakes)
set -u
mkdir -p gpurun_out
for v in exp0_trace exp2560_trace; do
  NERF_B200_LIB=$PWD/nerf-replication_b200/libnerf_b200_$v.so timeout 300 python profiles/trace_step.py 600 900 > gpurun_out/trace_$v.txt 2>&1
  echo "== $v"; head -3 gpurun_out/trace_$v.txt
done
