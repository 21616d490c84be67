#!/bin/bash
# round-2 GPU pass K: ncu --set full of the 16-warp fused-pool stem and of the kw-split layer1 kernel (no-residual and
# residual launch), raw + source pages exported on the box
mkdir -p gpurun_out
Q="--quick --no-cpu-baseline --no-e2e --sustained-seconds 0"
timeout 300 python bench.py --steps 2 --warmup 3 $Q > gpurun_out/plain_k.json 2> gpurun_out/plain_k.err; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_conv_stem|k_conv_flat3" -s 21 -c 3 -o /tmp/k_full python bench.py --steps 2 --warmup 3 $Q > gpurun_out/ncu_k.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/k_full.ncu-rep --page raw --csv > gpurun_out/r02_stem16_kw3_raw.csv 2> gpurun_out/export_k.err
ncu -i /tmp/k_full.ncu-rep --page source --csv > gpurun_out/r02_stem16_kw3_source.csv 2>> gpurun_out/export_k.err
ls -la gpurun_out/*.csv | tail -5
