#!/bin/bash
# round-2 GPU pass E: stain tests, timed default bench (as the driver runs it), reference arm, then the ncu evidence
mkdir -p gpurun_out
( time python -m pytest tests/test_stain.py tests/test_gpu_seg.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -8 ) > gpurun_out/pytest_e.log 2>&1
( time python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_e.json 2> gpurun_out/bench_e.err ) 2> gpurun_out/bench_e.time
( time python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_e.json 2> gpurun_out/bench_ref_e.err ) 2> gpurun_out/bench_ref_e.time
bash tools/gpu_profile_r2.sh > gpurun_out/profile.log 2>&1
echo done > gpurun_out/done.txt
