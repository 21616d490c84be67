#!/bin/bash
# A/B of the smem-ring pool kernel (InceptionV4) + ResNet34 with the residual L2 prefetch
mkdir -p gpurun_out
GPU_TEST_TIMEOUT=900 ./tools_gpu_tests.sh 2>&1 | grep "exit\|passed\|failed" | tail -20
timeout 600 python bench.py --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/ab_r34.json 2> gpurun_out/ab_r34.err; echo "r34 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/ab_r34.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'])"
grep "^op" gpurun_out/ab_r34.err | awk '{printf "%s:%s ", $2,$5}'; echo
for np in 0 1; do
  WSI_B200_NO_POOL_RING=$np timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/ab_iv4_noring$np.json 2> gpurun_out/ab_iv4_noring$np.err; echo "iv4 noring=$np rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/ab_iv4_noring$np.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['stages']['pool_head'])"
  grep "^op" gpurun_out/ab_iv4_noring$np.err | grep pool | awk '{printf "%s:%s ", $2,$5}'; echo
done
