#!/bin/bash
# round 2, calls J / K (NGPU = 2 or 4): parity checks of both set-up paths at this rank count; with BENCH=1 also the GPU arm at the default size
mkdir -p gpurun_out
N=${NGPU:-2}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 bash -c "$(declare -f run); N=$N run 29551 scripts/multi_gpu_local_check.py" > gpurun_out/r02j_local_check_n$N.raw 2>&1; echo "local check rc=$?"
grep "GPUs" gpurun_out/r02j_local_check_n$N.raw | tee gpurun_out/r02j_multi_gpu_local_check_n$N.log
timeout 600 bash -c "$(declare -f run); N=$N run 29552 scripts/multi_gpu_check.py" > gpurun_out/r02j_global_check_n$N.raw 2>&1; echo "global check rc=$?"
grep "GPUs" gpurun_out/r02j_global_check_n$N.raw | tee gpurun_out/r02j_multi_gpu_check_n$N.log
if [ -n "$BENCH" ]; then
  S=$SECONDS
  timeout 1200 bash -c "$(declare -f run); N=$N run 29553 bench.py --gpus $N --steps 5 --warmup 3" 2> gpurun_out/r02j_bench_n$N.err | grep '^{' > gpurun_out/r02j_bench_n$N.json; echo "gpu arm rc=${PIPESTATUS[0]} in $((SECONDS-S)) s"
  grep "\[bench\]" gpurun_out/r02j_bench_n$N.err; cut -c1-2200 gpurun_out/r02j_bench_n$N.json
fi
