#!/bin/bash
# round-2 GPU pass D: seg after the 4-pixels-per-thread change (tests, timing incl. cohort, launch list)
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_seg.py tests/test_cellpost.py tests/test_cli.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -15 ) > gpurun_out/pytest_seg_d.log 2>&1
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --parity-patches 2048 > gpurun_out/bench_d.json 2> gpurun_out/bench_d.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_d.csv \
  python bench.py --steps 2 --warmup 3 --quick --no-cpu-baseline --no-e2e --sustained-seconds 0 > gpurun_out/ncu_bench_d.log 2>&1
echo done > gpurun_out/done.txt
