#!/bin/bash
# First GPU call of round 2 (one GPU): everything written after round 1's GPU minutes ran out, in the order that matters.
#   gpurun --timeout 2400 -- 'bash scripts/gpu_round_r02.sh r02a'
# (1) the whole GPU suite (74 tests; test_zz_* have never run on a B200), (2) smoke, (3) the bench line,
# (4) the ncu launch list of the same command, (5) one --set full capture of the two sweeps, (6) setup with the batched device eigensolve.
TAG=${1:-r02a}
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 $OUT/${TAG}_smoke.log
PACK=/tmp/bench_112.ddmpack
timeout 900 python bench.py --pack $PACK > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"; tail -1 $OUT/${TAG}_bench.json | cut -c1-1200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --pack $PACK --steps 1 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_list.log 2>&1; echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:trisolve_(fwd|bwd)_kernel' -s 4 -c 2 -o $OUT/${TAG}_sweeps \
    python bench.py --pack $PACK --steps 1 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "set full exit $?"
DDM_SETUP_VERBOSE=1 timeout 900 python bench.py --eigensolver gpu --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_eiggpu.json 2> $OUT/${TAG}_bench_eiggpu.err
echo "device eigensolve exit $?"; grep -i "geneo\|eigen\|setup" $OUT/${TAG}_bench_eiggpu.err | tail -5; tail -1 $OUT/${TAG}_bench_eiggpu.json | cut -c1-600
# (7) optional, C5=1: config C5's per-GPU unit (232^3 nodes, 64 bricks of <= 62^3 overlapped nodes, about 100 GB of tiles), POU coarse space
if [ -n "$C5" ]; then
    free -g | head -2
    DDM_SETUP_VERBOSE=1 timeout 2400 python bench.py --workload q1cube --ncell 231 --nsub 64 --coarse pou --steps 2 --warmup 3 --no-cpu-baseline \
        > $OUT/${TAG}_bench_c5unit.json 2> $OUT/${TAG}_bench_c5unit.err
    echo "C5 unit exit $?"; tail -3 $OUT/${TAG}_bench_c5unit.err; tail -1 $OUT/${TAG}_bench_c5unit.json | cut -c1-900
fi
