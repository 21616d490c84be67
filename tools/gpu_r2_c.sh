#!/bin/bash
# round-2 GPU pass C: seg with one labelling + incremental merges (tests, timing, launch list)
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_seg.py tests/test_cellpost.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 ) > gpurun_out/pytest_seg.log 2>&1
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c.json 2> gpurun_out/bench_c.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_c.csv \
  python bench.py --steps 2 --warmup 3 --quick --no-cpu-baseline --no-e2e --sustained-seconds 0 > gpurun_out/ncu_bench_c.log 2>&1
echo done > gpurun_out/done.txt
