#!/bin/bash
# A/B of the round-2 conv kernel changes on one box: net tests, quick bench with per-op events for the default path,
# WSI_B200_KW3=1 (kw-split layer1 kernel) and WSI_B200_NO_DS_MERGE=1 (downsample convs as separate launches), then an
# ncu --set full capture of the fused-pool stem and (with WSI_B200_KW3=1) of the kw-split kernel.
mkdir -p gpurun_out
Q="--quick --no-cpu-baseline --no-e2e"
( time timeout 900 python -m pytest tests/test_gpu_net.py -x -q --tb=short -p no:cacheprovider 2>&1 | tail -60 ) > gpurun_out/pytest_conv_ab.log 2>&1
for v in default kw3 nomerge; do
  case $v in kw3) export WSI_B200_KW3=1;; nomerge) export WSI_B200_NO_DS_MERGE=1;; esac
  timeout 300 python bench.py --steps 20 --warmup 5 --per-op $Q --sustained-seconds 3 > gpurun_out/bench_ab_$v.json 2> gpurun_out/bench_ab_$v.err
  echo "bench rc=$?" >> gpurun_out/bench_ab_$v.err
  unset WSI_B200_KW3 WSI_B200_NO_DS_MERGE
done
WSI_B200_KW3=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_conv_stem|k_conv_flat3" -s 21 -c 3 -o /tmp/k_full python bench.py --steps 2 --warmup 3 $Q --sustained-seconds 0 > gpurun_out/ncu_conv_ab.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/k_full.ncu-rep --page raw --csv > gpurun_out/r02_stem16_kw3_raw.csv 2> gpurun_out/export_ab.err
ncu -i /tmp/k_full.ncu-rep --page source --csv > gpurun_out/r02_stem16_kw3_source.csv 2>> gpurun_out/export_ab.err
echo done > gpurun_out/done.txt
