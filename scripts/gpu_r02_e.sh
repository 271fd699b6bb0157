#!/bin/bash
# round 2, call E (one GPU): whole GPU suite on the new kernels (sliced-ELL SpMV, restrict, deterministic sweeps), both bench
# arms at the default (C5) size, deterministic sweeps timed at 160 cells, ncu traffic of the sweeps at the default size
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02e_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02e_suite.log
timeout 1500 python bench.py --impl reference --steps 3 --warmup 2 > gpurun_out/r02e_ref.json 2> gpurun_out/r02e_ref.err; echo "reference arm rc=$? ($SECONDS s so far)"
cut -c1-1800 gpurun_out/r02e_ref.json; tail -3 gpurun_out/r02e_ref.err
timeout 1500 python bench.py --steps 5 --warmup 3 > gpurun_out/r02e_bench.json 2> gpurun_out/r02e_bench.err; echo "gpu arm rc=$? ($SECONDS s so far)"
cut -c1-3500 gpurun_out/r02e_bench.json; tail -3 gpurun_out/r02e_bench.err
timeout 600 python bench.py --ncell 160 --deterministic --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r02e_det160.json 2> gpurun_out/r02e_det160.err; echo "det rc=$?"
timeout 600 python bench.py --ncell 160 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/r02e_nondet160.json 2> gpurun_out/r02e_nondet160.err; echo "nondet rc=$?"
python - <<'P'
import json
for f in ("r02e_det160", "r02e_nondet160"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, d["ms_per_step"], d["iterations"], d["roofline"]["kernels"]["trisolve_fwd"]["ms"], d["roofline"]["kernels"]["trisolve_bwd"]["ms"], d["timers_ms_per_call"])
    except Exception as e:
        print(f, "unreadable", e)
P
export DDM_PROFILE_REGION=1
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/r02e_plain.json 2> gpurun_out/r02e_plain.err &&
timeout 1500 ncu --profile-from-start off --set full --clock-control none -k regex:"trisolve_fwd|trisolve_bwd|restrict_kernel|spmv_sell" -c 4 -o gpurun_out/r02e_prof232 $CMD > gpurun_out/r02e_ncu.log 2>&1
echo "ncu rc=$? ($SECONDS s)"; tail -3 gpurun_out/r02e_ncu.log
