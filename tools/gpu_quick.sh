#!/bin/bash
# tests + both bench lines with per-op timings (no ncu)
mkdir -p gpurun_out
GPU_TEST_TIMEOUT=900 ./tools_gpu_tests.sh 2>&1 | grep -v "Warning\|warnings.warn" | tail -30
timeout 600 python bench.py --steps 30 --warmup 3 --per-op --no-cpu-baseline > gpurun_out/bench_r34.json 2> gpurun_out/bench_r34.err; echo "bench r34 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_r34.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages'],d['e2e'])"
grep "^op" gpurun_out/bench_r34.err | awk '{printf "%s:%s ", $2,$5}'; echo
timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline > gpurun_out/bench_iv4.json 2> gpurun_out/bench_iv4.err; echo "bench iv4 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_iv4.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages'],d['e2e'])"
