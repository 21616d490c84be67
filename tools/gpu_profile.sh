#!/bin/bash
# round-end evidence: ncu launch list of the bench command, one --set full capture of the conv kernels (ResNet34),
# DRAM/tensor metrics of every InceptionV4 conv.  Reports are exported to CSV ON THE BOX and deleted: gpurun only
# brings back 64 MiB.
mkdir -p gpurun_out
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,l1tex__m_xbar2l1tex_read_bytes.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_uniform.sum,smsp__cycles_active.avg,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,lts__t_bytes.sum"
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.json 2> gpurun_out/plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv -s 36 -c 36 -o /tmp/conv_full_r34 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_r34.log 2>&1; echo "ncu full r34 rc=$?"
ncu -i /tmp/conv_full_r34.ncu-rep --page raw --csv > gpurun_out/conv_full_r34_raw.csv 2> gpurun_out/export.err; echo "export raw rc=$?"
ncu -i /tmp/conv_full_r34.ncu-rep --page details --csv > gpurun_out/conv_full_r34_details.csv 2>> gpurun_out/export.err; echo "export details rc=$?"
for k in k_conv_flat k_conv_igemm k_conv_stem; do
  ncu -i /tmp/conv_full_r34.ncu-rep --page source --csv -k regex:$k -c 1 > gpurun_out/conv_full_r34_source_$k.csv 2>> gpurun_out/export.err; echo "export source $k rc=$?"
done
timeout 900 ncu --metrics $M --clock-control none -k regex:k_conv -s 133 -c 133 --csv --log-file gpurun_out/conv_metrics_iv4.csv python bench.py --model inceptionv4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_iv4.log 2>&1; echo "ncu iv4 rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_reduce|k_preprocess|k_pool3_tma|k_head" -c 8 -o /tmp/aux_full python bench.py --model inceptionv4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_aux.log 2>&1; echo "ncu aux rc=$?"
ncu -i /tmp/aux_full.ncu-rep --page raw --csv > gpurun_out/aux_full_raw.csv 2>> gpurun_out/export.err
cp /tmp/aux_full.ncu-rep gpurun_out/ 2>/dev/null
gzip -9 gpurun_out/conv_full_r34_source_*.csv
ls -la gpurun_out /tmp/*.ncu-rep
du -sh gpurun_out
