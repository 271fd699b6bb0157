#!/bin/bash
# One-GPU call: (1) ncu --set full of the secondary kernels on the bench workload, (2) the larger-subdomain Q1 workload.
TAG=${1:-r01f}
OUT=gpurun_out
mkdir -p $OUT
PACK=/tmp/bench_112.ddmpack
python bench.py --ncell 112 --pack $PACK --setup-only > $OUT/${TAG}_setup.log 2>&1
timeout 300 python scripts/profile_apply.py $PACK 2 > $OUT/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k 'regex:restrict_kernel|prolong_scale|spmv_csr|gather_sum|dense_gemv' -c 12 -o $OUT/${TAG}_secondary \
    python scripts/profile_apply.py $PACK 2 > $OUT/${TAG}_ncu.log 2>&1
tail -2 $OUT/${TAG}_ncu.log
DDM_SETUP_VERBOSE=1 timeout 1500 python bench.py --workload q1cube --ncell ${Q1N:-160} --nsub 64 --nev 16 > $OUT/${TAG}_bench_q1.json 2> $OUT/${TAG}_bench_q1.err
echo "q1 bench exit $?"
tail -3 $OUT/${TAG}_bench_q1.err
tail -1 $OUT/${TAG}_bench_q1.json | cut -c1-1500
