#!/bin/bash
# final state: the driver's invocations, then the ncu evidence
./tools/gpu_all.sh
timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline > gpurun_out/bench_iv4.json 2> gpurun_out/bench_iv4.err; echo "bench iv4 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_iv4.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['e2e'])"
./tools/gpu_profile.sh
