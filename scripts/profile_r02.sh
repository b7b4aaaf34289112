#!/bin/bash
# Round-2 profile capture (run on the GPU box through gpurun): launch lists + one full capture per dominant kernel,
# each after the same command has exited 0 without ncu; a compute-sanitizer pass over the small all-kernels script.
set -u
R=r02
mkdir -p gpurun_out
C3="python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu-baseline"
Q1="python bench.py --workload c3q1 --steps 5 --warmup 3 --no-cpu-baseline"
$C3 > gpurun_out/${R}_c3_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/${R}_c3_launches.csv $C3 > gpurun_out/${R}_c3_ncu1.log 2>&1
$C3 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k3_tf32 -s 0 -c 6 -o gpurun_out/${R}_k3_c3 $C3 > gpurun_out/${R}_c3_ncu2.log 2>&1
$Q1 > gpurun_out/${R}_q1_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${R}_q1_launches.csv $Q1 > gpurun_out/${R}_q1_ncu1.log 2>&1
$Q1 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:scan_l2 -s 4 -c 2 -o gpurun_out/${R}_k2_c3q1 $Q1 > gpurun_out/${R}_q1_ncu2.log 2>&1
# (compute-sanitizer is closed on this pool -- see profiles/r02_notes.md; the small all-kernels script runs plain)
python scripts/sanitize_small.py > gpurun_out/${R}_sanitize_plain.log 2>&1
ls -la gpurun_out | grep r02
