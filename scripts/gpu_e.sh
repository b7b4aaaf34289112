#!/bin/bash
mkdir -p gpurun_out
for args in "10000000 1024 100 384" "10000000 256 10 128" "4000000 1024 100 768"; do
  timeout 300 python scripts/probe_levels.py $args >> gpurun_out/e_probe.txt 2>&1
done
grep -E "N=|rror" gpurun_out/e_probe.txt
