#!/bin/bash
# e2e experiment: where do the 4 % between the resident and the host-buffer path go?
mkdir -p gpurun_out
Q="--quick --no-cpu-baseline --sustained-seconds 0"
python bench.py --steps 100 --warmup 5 $Q > gpurun_out/e2e_k100.json 2> gpurun_out/e2e_k100.err
WSI_B200_HP_NO_H2D=1 python bench.py --steps 100 --warmup 5 $Q > gpurun_out/e2e_k100_noh2d.json 2> gpurun_out/e2e_k100_noh2d.err
python bench.py --steps 20 --warmup 5 $Q > gpurun_out/e2e_k20.json 2> gpurun_out/e2e_k20.err
( time python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "config1 or resnet34-2" 2>&1 | tail -5 ) > gpurun_out/pytest_f.log 2>&1
echo done > gpurun_out/done.txt
