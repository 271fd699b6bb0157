#!/bin/bash
# round 2, call C: rank-local path check (world 1), ncu launch list + full capture of the solve-phase kernels at the bench's
# default size, then the bench at config C5's own brick size (232 cells per axis: 58 owned cells, 63^3-node overlapping bricks)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py tests/test_zz_gpu_parity_report.py -q -m gpu -k "local or C3 or C4" > gpurun_out/r02c_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r02c_tests.log
export DDM_PROFILE_REGION=1
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/r02c_plain.json 2> gpurun_out/r02c_plain.err &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r02c_launches.csv $CMD > gpurun_out/r02c_ncu1.log 2>&1
echo "ncu launch list rc=$?"
$CMD > gpurun_out/r02c_plain2.json 2> gpurun_out/r02c_plain2.err &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"trisolve_fwd_kernel|trisolve_bwd_kernel|restrict_kernel|spmv_csr_kernel|prolong_scale_kernel|gather_sum_kernel" -c 6 -o gpurun_out/r02c_prof $CMD > gpurun_out/r02c_ncu2.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/r02c_ncu2.log
unset DDM_PROFILE_REGION
ls -la gpurun_out/ | head -30
DDM_SETUP_VERBOSE=1 timeout 1500 python bench.py --ncell 232 --steps 3 --warmup 2 > gpurun_out/r02c_bench232.json 2> gpurun_out/r02c_bench232.err; echo "bench232 rc=$?"
tail -20 gpurun_out/r02c_bench232.err; cut -c1-2500 gpurun_out/r02c_bench232.json
nvidia-smi --query-gpu=memory.used --format=csv
