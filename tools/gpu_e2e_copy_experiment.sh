#!/bin/bash
# N=8 experiment: is the e2e droop the host->device copies?  (RGB vs RGBA host buffers are both in the bench line)
N=8
mkdir -p gpurun_out
Q="--quick --no-cpu-baseline --sustained-seconds 0"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
$TR bench.py --gpus $N --steps 20 --warmup 5 $Q > gpurun_out/n8_e2e.json 2> gpurun_out/n8_e2e.err
WSI_B200_HP_NO_H2D=1 $TR bench.py --gpus $N --steps 20 --warmup 5 $Q > gpurun_out/n8_e2e_noh2d.json 2> gpurun_out/n8_e2e_noh2d.err
echo done > gpurun_out/done.txt
