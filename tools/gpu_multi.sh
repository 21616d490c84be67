#!/bin/bash
# round-2 multi-GPU pass (run with gpurun --gpus N): the 2-rank CLI test and the bench under torchrun
N=${1:-2}
mkdir -p gpurun_out
( time python -m pytest tests/test_cli.py -m gpu -q -x -p no:cacheprovider -k two_ranks 2>&1 | tail -15 ) > gpurun_out/pytest_multi.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
echo done > gpurun_out/done_multi.txt
