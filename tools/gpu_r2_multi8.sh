#!/bin/bash
# 8-GPU check of the bench line exactly as the driver launches it
N=${1:-8}
mkdir -p gpurun_out
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err ) 2> gpurun_out/bench_${N}gpu.time
echo "rc=$?" >> gpurun_out/bench_${N}gpu.err
nvidia-smi topo -m > gpurun_out/topo_${N}.txt 2>&1
echo done > gpurun_out/done_multi.txt
