#!/bin/bash
# round 2, call F (N = 2 or 4 GPUs): rank-local path against the oracle at this rank count (small case), the global-path halo
# check, then both bench arms at 160 cells per GPU axis under torchrun
mkdir -p gpurun_out
N=${NGPU:-4}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 bash -c "$(declare -f run); N=$N run 29541 scripts/multi_gpu_local_check.py" > gpurun_out/r02f_local_check_n$N.log 2>&1; echo "local check rc=$?"
grep "GPUs" gpurun_out/r02f_local_check_n$N.log
timeout 600 bash -c "$(declare -f run); N=$N run 29542 scripts/multi_gpu_check.py" > gpurun_out/r02f_global_check_n$N.log 2>&1; echo "global check rc=$?"
grep "GPUs" gpurun_out/r02f_global_check_n$N.log
timeout 900 bash -c "$(declare -f run); N=$N run 29543 bench.py --impl reference --gpus $N --ncell ${NCELL:-160} --steps 2 --warmup 1" > gpurun_out/r02f_ref_n$N.json 2> gpurun_out/r02f_ref_n$N.err; echo "reference arm rc=$?"
cut -c1-1500 gpurun_out/r02f_ref_n$N.json; tail -2 gpurun_out/r02f_ref_n$N.err | cut -c1-300
timeout 900 bash -c "$(declare -f run); N=$N run 29544 bench.py --gpus $N --ncell ${NCELL:-160} --steps 3 --warmup 2" > gpurun_out/r02f_bench_n$N.json 2> gpurun_out/r02f_bench_n$N.err; echo "gpu arm rc=$?"
cut -c1-2500 gpurun_out/r02f_bench_n$N.json; grep "\[bench\]" gpurun_out/r02f_bench_n$N.err
