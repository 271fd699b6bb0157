#!/bin/bash
# round 2, call D (8 GPUs): box resources, rank-local path against the oracle at world 8 (small case), then the bench at config
# C5's own size: 464^3 cells = 99.9 M DOFs, 8 x 8 x 8 = 512 brick subdomains
mkdir -p gpurun_out
{ nproc; free -g | head -2; nvidia-smi --query-gpu=index,name,memory.total --format=csv; python -c "import os;print('affinity', len(os.sched_getaffinity(0)))"; } > gpurun_out/r02d_box.txt 2>&1
cat gpurun_out/r02d_box.txt
N=${NGPU:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 scripts/multi_gpu_local_check.py > gpurun_out/r02d_local_check_n$N.log 2>&1; echo "local check rc=$?"
grep "GPUs" gpurun_out/r02d_local_check_n$N.log
DDM_SETUP_VERBOSE=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --ncell ${NCELL:-232} --steps 3 --warmup 2 > gpurun_out/r02d_bench_n$N.json 2> gpurun_out/r02d_bench_n$N.err; echo "bench rc=$?"
grep "setup r0\|\[bench\]" gpurun_out/r02d_bench_n$N.err | tail -12; tail -3 gpurun_out/r02d_bench_n$N.err | cut -c1-400
cut -c1-3000 gpurun_out/r02d_bench_n$N.json
free -g | head -2
