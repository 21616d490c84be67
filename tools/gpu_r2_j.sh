#!/bin/bash
# round-2 GPU pass J: conv kernel changes (16-warp fused-pool stem epilogue, kw-split layer1 kernel): net tests + quick bench A/B
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_net.py -x -q --tb=short -p no:cacheprovider 2>&1 | tail -60 ) > gpurun_out/pytest_j.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --per-op --quick --no-cpu-baseline --no-e2e --sustained-seconds 3 > gpurun_out/bench_j.json 2> gpurun_out/bench_j.err
echo "bench rc=$?" >> gpurun_out/bench_j.err
WSI_B200_NO_KW3=1 timeout 300 python bench.py --steps 20 --warmup 5 --per-op --quick --no-cpu-baseline --no-e2e --sustained-seconds 3 > gpurun_out/bench_j_nokw3.json 2> gpurun_out/bench_j_nokw3.err
echo "bench rc=$?" >> gpurun_out/bench_j_nokw3.err
echo done > gpurun_out/done.txt
