#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/diag_net.py > gpurun_out/diag_net.log 2>&1; echo "diag exit $?"
grep -v Warning gpurun_out/diag_net.log | grep -v warnings.warn | head -150
GPU_TEST_TIMEOUT=600 ./tools_gpu_tests.sh 2>&1 | grep -v "Warning\|warnings.warn" | tail -60
