#!/bin/bash
# the driver's own invocations: one pytest process over all GPU tests, smoke(), default bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu_all.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 30 --warmup 3 --per-op > gpurun_out/bench_r34.json 2> gpurun_out/bench_r34.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_r34.json'));print(d['value'],d['e2e']['value'],d['stages']['density_grid'],d['stages']['thumbnail']['ms_per_slide'],d['cpu_baseline'])"
