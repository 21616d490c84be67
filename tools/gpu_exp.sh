#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_preproc.py tests/test_image_cache.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
for m in resnet34 inceptionv4; do
timeout 300 python bench.py --model $m --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/exp_$m.json 2> gpurun_out/exp_$m.err
python -c "import json;d=json.load(open('gpurun_out/exp_$m.json'));print('$m',d['value'],d['ms_per_step'],d['stages']['preprocess'])"
done
