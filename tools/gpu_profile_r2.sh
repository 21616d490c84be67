#!/bin/bash
# round-2 ncu evidence: --set full capture of the ResNet34 conv kernels (one forward), DRAM / tensor metrics of every
# InceptionV4 conv, --set full of the new seg and cell-path kernels.  Reports are exported to CSV ON THE BOX (gpurun only
# brings back 64 MiB).
mkdir -p gpurun_out
Q="--quick --no-cpu-baseline --no-e2e --sustained-seconds 0"
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,l1tex__m_xbar2l1tex_read_bytes.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,lts__t_bytes.sum"
timeout 600 python bench.py --steps 2 --warmup 3 $Q > gpurun_out/plain.json 2> gpurun_out/plain.err; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv -s 36 -c 36 -o /tmp/conv_full_r34 python bench.py --steps 2 --warmup 3 $Q > gpurun_out/ncu_full_r34.log 2>&1; echo "ncu full r34 rc=$?"
ncu -i /tmp/conv_full_r34.ncu-rep --page raw --csv > gpurun_out/r02_conv_full_r34_raw.csv 2> gpurun_out/export.err; echo "export raw rc=$?"
timeout 900 ncu --metrics $M --clock-control none -k regex:k_conv -s 119 -c 119 --csv --log-file gpurun_out/r02_conv_metrics_iv4.csv python bench.py --model inceptionv4 --steps 2 --warmup 3 $Q > gpurun_out/ncu_iv4.log 2>&1; echo "ncu iv4 rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_sat_median|k_closing|k_ccl_|k_outer_|k_grid_" -s 16 -c 16 -o /tmp/seg_full python bench.py --steps 2 --warmup 3 $Q > gpurun_out/ncu_seg.log 2>&1; echo "ncu seg rc=$?"
ncu -i /tmp/seg_full.ncu-rep --page raw --csv > gpurun_out/r02_seg_full_raw.csv 2>> gpurun_out/export.err
timeout 600 ncu --set full --clock-control none -k regex:"k_cp_" -c 24 -o /tmp/cp_full python tools/cellpost_bench.py > gpurun_out/ncu_cp.log 2>&1; echo "ncu cellpost rc=$?"
ncu -i /tmp/cp_full.ncu-rep --page raw --csv > gpurun_out/r02_cellpost_full_raw.csv 2>> gpurun_out/export.err
ls -la gpurun_out /tmp/*.ncu-rep; du -sh gpurun_out
