#!/bin/bash
# final-state verification pass: full GPU suite, smoke, default bench + reference arm, launch list, ncu --set full of
# the 33 ResNet34 conv launches of one forward (traffic JSON + summary)
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider 2>&1 | tail -60 ) > gpurun_out/pytest_m.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_m.log 2>&1
( time python bench.py --gpus 1 --steps 20 --warmup 5 --per-op > gpurun_out/bench_m.json 2> gpurun_out/bench_m.err ) 2> gpurun_out/bench_m.time
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_m.json 2> gpurun_out/bench_ref_m.err
Q="--quick --no-cpu-baseline --no-e2e --sustained-seconds 0"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_final.csv python bench.py --steps 2 --warmup 3 $Q > gpurun_out/ncu_launches_m.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv -s 33 -c 33 -o /tmp/conv_full_m python bench.py --steps 2 --warmup 3 $Q > gpurun_out/ncu_full_m.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/conv_full_m.ncu-rep --page raw --csv > gpurun_out/r02_conv_full_final_raw.csv 2> gpurun_out/export_m.err
echo done > gpurun_out/done.txt
