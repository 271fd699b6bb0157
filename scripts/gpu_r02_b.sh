#!/bin/bash
# round 2, call B: parity report at config sizes, generic preconditioner list, bench on the C5 family (N = 1, default size)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests/test_zz_gpu_parity_report.py tests/test_ref_krylov.py -q -m gpu > gpurun_out/r02b_parity.log 2>&1; echo "parity rc=$?"
grep "PARITY\|passed\|failed" gpurun_out/r02b_parity.log | cut -c1-600
DDM_SETUP_VERBOSE=1 timeout 1500 python bench.py --steps 5 --warmup 3 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench rc=$?"
tail -25 gpurun_out/r02b_bench.err; cat gpurun_out/r02b_bench.json | cut -c1-3000
