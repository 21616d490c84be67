#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pool -s 2 -c 2 -o gpurun_out/pool_r34 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_pool_r34.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pool -s 18 -c 5 -o gpurun_out/pool_iv4 python bench.py --model inceptionv4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_pool_iv4.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_preprocess -s 2 -c 1 -o gpurun_out/preproc_r34 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_pre_r34.log 2>&1; echo "rc=$?"
ls -la gpurun_out/*.ncu-rep
