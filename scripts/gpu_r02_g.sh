#!/bin/bash
# round 2, call G (one GPU): whole GPU suite (no -x: every failure at once), smoke(), a short default bench for the coloured Galerkin product
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -q -m gpu > gpurun_out/r02g_suite.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r02g_suite.log | cut -c1-250
timeout 600 python __graft_entry__.py smoke > gpurun_out/r02g_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02g_smoke.log
timeout 900 python bench.py --ncell 160 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r02g_bench160.json 2> gpurun_out/r02g_bench160.err; echo "bench160 rc=$?"
cat gpurun_out/r02g_bench160.err; cut -c1-1200 gpurun_out/r02g_bench160.json
