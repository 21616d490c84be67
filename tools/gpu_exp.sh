#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py tests/test_image_cache.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
run() {
  tag=$1; shift
  env "$@" timeout 300 python bench.py --model ${MODEL:-resnet34} --steps 30 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/exp_$tag.json 2> gpurun_out/exp_$tag.err
  python -c "import json;d=json.load(open('gpurun_out/exp_$tag.json'));print('$tag',d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages']['conv'])"
}
run pdl X=1
run nopdl WSI_B200_NO_PDL=1
run pdl2 X=1
run nopdl2 WSI_B200_NO_PDL=1
MODEL=inceptionv4 run iv4_pdl X=1
MODEL=inceptionv4 run iv4_nopdl WSI_B200_NO_PDL=1
