#!/bin/bash
# Two-GPU call: GPU parity tests (incl. the torchrun world=2 check) and the bench line at N=2.
#   gpurun --gpus 2 --timeout 1500 -- 'bash scripts/gpu_round2.sh [tag] [ngpus]'
TAG=${1:-r01}
N=${2:-2}
OUT=gpurun_out
mkdir -p $OUT
nproc; free -g | head -2; nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu_n$N.log 2>&1
echo "pytest exit $?" >> $OUT/${TAG}_pytest_gpu_n$N.log
tail -4 $OUT/${TAG}_pytest_gpu_n$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N ${NCELL:+--ncell $NCELL} > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err
echo "bench exit $?"
tail -3 $OUT/${TAG}_bench_n$N.err
tail -1 $OUT/${TAG}_bench_n$N.json | cut -c1-1500
