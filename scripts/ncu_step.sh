#!/bin/bash
# One un-graphed training step under ncu --set full (with source).  Run through gpurun on ONE GPU; ncu runs only
# after the same command exited 0 without it.  profile_step.py brackets its last step with
# cudaProfilerStart/Stop, so --profile-from-start off captures exactly one step.  The raw page is exported on the
# box and the report is compressed so that gpurun_out/ stays under the 64 MiB that travels back.
set -e
mkdir -p gpurun_out
rm -f gpurun_out/*.ncu-rep gpurun_out/*.ncu-rep.gz
CMD="python scripts/profile_step.py 3 64"
$CMD > gpurun_out/plain_step.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -f -o gpurun_out/step_full_r2 $CMD \
    > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/step_full_r2.ncu-rep --page raw --csv > gpurun_out/step_full_r2_raw.csv
gzip -9 gpurun_out/step_full_r2.ncu-rep
SZ=$(stat -c %s gpurun_out/step_full_r2.ncu-rep.gz)
echo "compressed report: $SZ bytes"
if [ "$SZ" -gt 50000000 ]; then rm gpurun_out/step_full_r2.ncu-rep.gz; echo "report dropped (too large)"; fi
du -sh gpurun_out
