#!/bin/bash
# round 2, call A: host resources of the box, device factorisation tests, the existing parity suite on the refactored upload path
mkdir -p gpurun_out
{ nproc; free -g; lscpu | grep -i "model name\|socket\|thread\|^CPU(s)"; nvidia-smi --query-gpu=name,memory.total --format=csv;
  python -c "import os;print('affinity', len(os.sched_getaffinity(0)))"; cat /sys/fs/cgroup/memory.max /sys/fs/cgroup/cpu.max 2>/dev/null; df -h /tmp | tail -1; } > gpurun_out/r02a_box.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_factorize.py -x -q -m gpu > gpurun_out/r02a_factorize.log 2>&1; echo "factorize rc=$?"
tail -15 gpurun_out/r02a_factorize.log
timeout 1500 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_factorize.py > gpurun_out/r02a_suite.log 2>&1; echo "suite rc=$?"
tail -8 gpurun_out/r02a_suite.log
cat gpurun_out/r02a_box.txt
