#!/bin/bash
# Runs on the GPU box (under gpurun): plain run first, then the ncu launch list and one
# --set full capture of the hot kernels, as /opt/skills/guides/B200_PROFILING.md prescribes.
# Outputs land in gpurun_out/; summaries are copied into profiles/ by hand afterwards.
set -u
TAG=${1:-r1}
# --partition streams: ncu fails with "Failed to prepare kernel for profiling" on launches into green contexts
ARGS="--steps 2 --warmup 3 --no-cpu --vgpu-gib 4 --spill-chunks 64 --partition streams --config5-gib 1 --e2e-chunks 64"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu_list.log 2>&1
echo "launch list rc=$?"
python scripts/prof_kernels.py 4 > gpurun_out/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_memtest|k_xfer' -c 13 \
    -o gpurun_out/${TAG}_prof python scripts/prof_kernels.py 4 > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/
