#!/bin/bash
# round 2, call I (one GPU): reference arm with the driver's flags at the default (C5) size; ncu launch list of the default GPU-arm command
mkdir -p gpurun_out
S=$SECONDS
timeout 1500 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 2> gpurun_out/r02i_ref.err | grep '^{' > gpurun_out/r02i_ref.json; echo "reference arm rc=${PIPESTATUS[0]} in $((SECONDS-S)) s"
cut -c1-1500 gpurun_out/r02i_ref.json; tail -2 gpurun_out/r02i_ref.err | cut -c1-300
export DDM_PROFILE_REGION=1
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
S=$SECONDS
$CMD > gpurun_out/r02i_plain.out 2> gpurun_out/r02i_plain.err &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/r02i_launches_q1cube232.csv $CMD > gpurun_out/r02i_ncu.log 2>&1
echo "ncu launch list rc=$? in $((SECONDS-S)) s"
grep '^{' gpurun_out/r02i_plain.out | cut -c1-600
wc -l gpurun_out/r02i_launches_q1cube232.csv
