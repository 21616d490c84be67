#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_net.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -3
run() {
  tag=$1; shift
  env "$@" timeout 300 python bench.py --model ${MODEL:-resnet34} --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/exp_$tag.json 2> gpurun_out/exp_$tag.err
  python -c "import json;d=json.load(open('gpurun_out/exp_$tag.json'));print('$tag',d['value'],d['ms_per_step'],d['roofline']['achieved'])"
  grep "^op" gpurun_out/exp_$tag.err | head -20 | awk '{printf "%s:%s ", $2,$5}'; echo
}
run ressmem X=1
run resregs WSI_B200_NO_RES_SMEM=1
