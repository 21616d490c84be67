#!/bin/bash
# round-2 GPU pass L: downsample convs merged into the block's first conv (centre-tap-only N tiles): net tests + bench A/B
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_net.py -x -q --tb=short -p no:cacheprovider 2>&1 | tail -60 ) > gpurun_out/pytest_l.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --per-op --quick --no-cpu-baseline --no-e2e --sustained-seconds 3 > gpurun_out/bench_l.json 2> gpurun_out/bench_l.err
echo "bench rc=$?" >> gpurun_out/bench_l.err
WSI_B200_NO_DS_MERGE=1 timeout 300 python bench.py --steps 20 --warmup 5 --per-op --quick --no-cpu-baseline --no-e2e --sustained-seconds 3 > gpurun_out/bench_l_nomerge.json 2> gpurun_out/bench_l_nomerge.err
echo "bench rc=$?" >> gpurun_out/bench_l_nomerge.err
timeout 300 python bench.py --model inceptionv4 --steps 10 --warmup 3 --quick --no-cpu-baseline --no-e2e --sustained-seconds 0 > gpurun_out/bench_l_iv4.json 2> gpurun_out/bench_l_iv4.err
echo "bench rc=$?" >> gpurun_out/bench_l_iv4.err
echo done > gpurun_out/done.txt
