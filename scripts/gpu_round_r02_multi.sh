#!/bin/bash
# Multi-GPU call of round 2: the refined-unit-cube weak-scaling family (CPU oracle: 17 iterations at the N = 1 and N = 2 shapes) and the
# parity script.   gpurun --gpus N --timeout 2400 -- 'bash scripts/gpu_round_r02_multi.sh N r02b'
N=${1:-2}; TAG=${2:-r02b}
OUT=gpurun_out; mkdir -p $OUT
nproc; free -g | head -2
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k "$N" > $OUT/${TAG}_pytest_n$N.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest_n$N.log
for extra in "" "--strong" "--workload q1cube --ncell 96"; do
    name=$(echo "n${N}${extra}" | tr -d ' -')
    timeout 1800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 \
        bench.py --gpus $N $extra --no-cpu-baseline > $OUT/${TAG}_bench_$name.json 2> $OUT/${TAG}_bench_$name.err
    echo "bench $name exit $?"; tail -1 $OUT/${TAG}_bench_$name.json | cut -c1-900
done
