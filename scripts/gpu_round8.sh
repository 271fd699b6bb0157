#!/bin/bash
# Eight-GPU call: the world=8 parity check and one bench line (NCELL may shrink the per-GPU problem to save box time).
TAG=${1:-r01n}
OUT=gpurun_out; mkdir -p $OUT
nproc; free -g | head -2
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k "8" > $OUT/${TAG}_pytest_n8.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest_n8.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 \
    bench.py --gpus 8 ${NCELL:+--ncell $NCELL} > $OUT/${TAG}_bench_n8.json 2> $OUT/${TAG}_bench_n8.err
echo "bench exit $?"; grep "bench\]" $OUT/${TAG}_bench_n8.err; tail -1 $OUT/${TAG}_bench_n8.json | cut -c1-900
free -g | head -2
