#!/bin/bash
# round 2, call H (one GPU): ddm_dot_noalloc -- its test alone in a fresh process, then the whole GPU suite with -x (the test is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02h_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r02h_noalloc_alone.log | cut -c1-300
timeout 1700 python -m pytest tests -x -q -m gpu > gpurun_out/r02h_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02h_suite.log | cut -c1-300
grep -c "test_zz_no_alloc" gpurun_out/r02h_suite.log
timeout 1700 python -m pytest tests -q -m gpu -v 2>/dev/null | grep "no_alloc" | cut -c1-160
