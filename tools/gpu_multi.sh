#!/bin/bash
# what the driver does for the scaling bench: torchrun, one rank per GPU
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 3 > gpurun_out/bench_r34_n$N.json 2> gpurun_out/bench_r34_n$N.err; echo "rc=$?"
cat gpurun_out/bench_r34_n$N.json | cut -c1-3000
tail -5 gpurun_out/bench_r34_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
cat gpurun_out/bench_ref_n$N.json | cut -c1-600
