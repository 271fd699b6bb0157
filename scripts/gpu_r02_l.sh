#!/bin/bash
# round 2, call L (8 GPUs): the final code at 8 ranks -- rank-local parity check, then the GPU arm at 160 cells per GPU axis (320^3 cells)
mkdir -p gpurun_out
N=8
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 400 bash -c "$(declare -f run); N=$N run 29561 scripts/multi_gpu_local_check.py" > gpurun_out/r02l_local_check_n8.raw 2>&1; echo "local check rc=$?"
grep "GPUs" gpurun_out/r02l_local_check_n8.raw | tee gpurun_out/r02l_multi_gpu_local_check_n8.log
S=$SECONDS
timeout 600 bash -c "$(declare -f run); N=$N run 29563 bench.py --gpus $N --ncell 160 --steps 5 --warmup 3" 2> gpurun_out/r02l_bench_n8.err | grep '^{' > gpurun_out/r02l_bench_n8.json; echo "gpu arm rc=${PIPESTATUS[0]} in $((SECONDS-S)) s"
grep "\[bench\]" gpurun_out/r02l_bench_n8.err; cut -c1-1800 gpurun_out/r02l_bench_n8.json
