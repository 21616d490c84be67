#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pool -s 18 -c 4 -o gpurun_out/pool_iv4 python bench.py --model inceptionv4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_pool_iv4.log 2>&1; echo "rc=$?"
ls -la gpurun_out/*.ncu-rep
