#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests/test_image_cache.py tests/test_gpu_preproc.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -8 ) > gpurun_out/pytest_i.log 2>&1
python bench.py --steps 20 --warmup 5 --quick --no-cpu-baseline --sustained-seconds 0 > gpurun_out/bench_i.json 2> gpurun_out/bench_i.err
echo done > gpurun_out/done.txt
