#!/bin/bash
# final state of the round: driver invocations, InceptionV4 line, refreshed launch list + InceptionV4 conv metrics
./tools/gpu_all.sh
timeout 600 python bench.py --model inceptionv4 --steps 10 --warmup 3 --per-op --no-cpu-baseline > gpurun_out/bench_iv4.json 2> gpurun_out/bench_iv4.err; echo "bench iv4 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_iv4.json'));print(d['value'],d['ms_per_step'],d['roofline']['achieved'],d['e2e'])"
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,l1tex__m_xbar2l1tex_read_bytes.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics $M --clock-control none -k regex:k_conv -s 119 -c 119 --csv --log-file gpurun_out/conv_metrics_iv4.csv python bench.py --model inceptionv4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_iv4.log 2>&1; echo "ncu iv4 rc=$?"
