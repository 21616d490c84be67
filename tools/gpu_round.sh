#!/bin/bash
# One GPU visit: parity tests, bench lines, per-op timings, ncu launch list, one --set full capture.
mkdir -p gpurun_out
GPU_TEST_TIMEOUT=900 ./tools_gpu_tests.sh 2>&1 | grep -v "Warning\|warnings.warn" | tail -30
timeout 600 python bench.py --steps 30 --warmup 3 --per-op > gpurun_out/bench_r34.json 2> gpurun_out/bench_r34.err; echo "bench r34 rc=$?"
cat gpurun_out/bench_r34.json
timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline > gpurun_out/bench_iv4.json 2> gpurun_out/bench_iv4.err; echo "bench iv4 rc=$?"
cat gpurun_out/bench_iv4.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv -s 40 -c 36 -o gpurun_out/conv_full python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
