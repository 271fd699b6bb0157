#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the ncu launch list and one full capture of the sweeps.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh [tag]'
TAG=${1:-r01}
NCELL=${NCELL:-112}
OUT=gpurun_out
mkdir -p $OUT
PACK=/tmp/bench_$NCELL.ddmpack
set -x
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu.log
tail -3 $OUT/${TAG}_pytest_gpu.log
timeout 900 python bench.py --ncell $NCELL --pack $PACK > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"
tail -1 $OUT/${TAG}_bench.json | cut -c1-600
if [ -z "$SKIP_NCU" ]; then
timeout 300 python scripts/profile_apply.py $PACK 3 pcg > $OUT/${TAG}_profile_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python scripts/profile_apply.py $PACK 3 pcg > $OUT/${TAG}_ncu_launches.log 2>&1
timeout 300 python scripts/profile_apply.py $PACK 2 > $OUT/${TAG}_profile_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:trisolve -c 4 -o $OUT/${TAG}_trisolve \
    python scripts/profile_apply.py $PACK 2 > $OUT/${TAG}_ncu_full.log 2>&1
fi
ls -la $OUT
