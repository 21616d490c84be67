#!/bin/bash
# Run each GPU test file in its own process under a hard timeout (a hung kernel cannot stall the box).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for f in tests/test_gpu_seg.py tests/test_gpu_preproc.py tests/test_thumbnail.py tests/test_density.py tests/test_image_cache.py tests/test_tilefuse.py tests/test_gpu_net.py; do
  n=$(basename $f .py)
  timeout -k 10 ${GPU_TEST_TIMEOUT:-600} python -m pytest $f -m gpu -q -x --no-header -p no:cacheprovider ${PYTEST_EXTRA} > gpurun_out/$n.log 2>&1
  echo "$n exit $?" | tee -a gpurun_out/summary.txt
  tail -n 40 gpurun_out/$n.log
done
