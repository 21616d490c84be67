#!/bin/bash
# net tests + InceptionV4 per-op bench (pool / conv experiments)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_net.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/iv4.json 2> gpurun_out/iv4.err; echo "iv4 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/iv4.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages']['pool_head'], d['stages']['conv'])"
grep "^op" gpurun_out/iv4.err | grep pool | awk '{printf "%s:%s ", $2,$5}'; echo
grep "^op" gpurun_out/iv4.err | head -14 | awk '{printf "%s:%s ", $2,$5}'; echo
