#!/bin/bash
# A/B of the flat-kernel CTA pairs: tests, then per-op bench with and without pairs on the same box
mkdir -p gpurun_out
GPU_TEST_TIMEOUT=900 ./tools_gpu_tests.sh 2>&1 | grep -v "Warning\|warnings.warn" | tail -30
for m in resnet34 inceptionv4; do
  for np in 0 1; do
    WSI_B200_NO_FLAT_PAIR=$np timeout 600 python bench.py --model $m --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/ab_${m}_nopair$np.json 2> gpurun_out/ab_${m}_nopair$np.err; echo "$m nopair=$np rc=$?"
    python -c "import json;d=json.load(open('gpurun_out/ab_${m}_nopair$np.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages'])"
    if [ $m = resnet34 ]; then grep "^op" gpurun_out/ab_${m}_nopair$np.err | awk '{printf "%s:%s ", $2,$5}'; echo; fi
  done
done
