#!/bin/bash
# ncu --set full of the kernels matching $1 (regex on the function name) in the last of three un-graphed steps;
# exports the raw page as gpurun_out/$2_raw.csv and keeps the compressed report.  One GPU, after a plain run.
set -e
mkdir -p gpurun_out
CMD="python scripts/profile_step.py 3 64"
$CMD > gpurun_out/plain_step.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$1" -f -o gpurun_out/$2 $CMD \
    > gpurun_out/ncu_$2.log 2>&1
ncu -i gpurun_out/$2.ncu-rep --page raw --csv > gpurun_out/$2_raw.csv
gzip -9f gpurun_out/$2.ncu-rep
ls -la gpurun_out/$2*
