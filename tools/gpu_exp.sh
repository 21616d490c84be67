#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_image_cache.py -x -q -m gpu -p no:cacheprovider 2>&1 | grep -v Warn | tail -25
run() {
  tag=$1; shift
  env "$@" timeout 300 python bench.py --model ${MODEL:-resnet34} --steps 10 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/exp_$tag.json 2> gpurun_out/exp_$tag.err
  python -c "import json;d=json.load(open('gpurun_out/exp_$tag.json'));print('$tag',d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages']['conv'],d['stages']['pool_head'])"
  grep "^op" gpurun_out/exp_$tag.err | sed -n 12,22p | awk '{printf "%s:%s:%s ", $2,$3,$5}'; echo
}
MODEL=inceptionv4 run iv4_commute X=1
MODEL=inceptionv4 run iv4_nocommute WSI_B200_NO_POOL_COMMUTE=1
MODEL=inceptionv4 run iv4_nosib WSI_B200_NO_SIBLING_MERGE=1
