#!/bin/bash
# round-2 GPU pass A: GPU tests (incl. the >= 10^4-patch parity harness), bench, reference arm, seg A/B, launch list
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
python -c "import os; print(len(os.sched_getaffinity(0)))" > gpurun_out/cores.txt
( time python -m pytest tests -m gpu -q -p no:cacheprovider 2>&1 | tail -60 ) > gpurun_out/pytest_gpu.log 2>&1
python bench.py --steps 20 --warmup 5 --per-op > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?" >> gpurun_out/bench.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
WSI_B200_SEG_LEGACY=1 python bench.py --steps 10 --warmup 3 --quick --no-cpu-baseline --no-e2e --sustained-seconds 0 > gpurun_out/bench_seg_legacy.json 2> gpurun_out/bench_seg_legacy.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --quick --no-cpu-baseline --no-e2e --sustained-seconds 0 > gpurun_out/ncu_bench.log 2>&1
echo done > gpurun_out/done.txt
