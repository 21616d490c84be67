#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_image_cache.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
run() {
  tag=$1; shift
  env "$@" timeout 300 python bench.py --model ${MODEL:-resnet34} --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/exp_$tag.json 2> gpurun_out/exp_$tag.err
  python -c "import json;d=json.load(open('gpurun_out/exp_$tag.json'));print('$tag',d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages']['conv'])"
  grep "^op" gpurun_out/exp_$tag.err | head -${NOPS:-20} | awk '{printf "%s:%s ", $2,$5}'; echo
}
NOPS=32 run s2pair X=1
NOPS=32 run nos2pair WSI_B200_NO_S2PAIR=1
MODEL=inceptionv4 NOPS=12 run iv4_s2pair X=1
MODEL=inceptionv4 NOPS=12 run iv4_nos2pair WSI_B200_NO_S2PAIR=1
