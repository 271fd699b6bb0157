#!/bin/bash
# Timing experiments on the sweeps with one cached setup (apply timers only; no PCG: results may be wrong by design).
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT
PACK=/tmp/bench_ab.ddmpack
python bench.py --pack $PACK --setup-only > $OUT/${TAG}_setup.log 2>&1
for envs in "$@"; do
  echo "== $envs"
  env $envs python scripts/profile_apply.py $PACK 20 2>&1 | grep -v launches
done
