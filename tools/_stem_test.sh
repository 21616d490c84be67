mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_gpu_net.py -x -q --tb=short -p no:cacheprovider -k "resnet34 or stem or downsample" 2>&1 | tail -30 ) > gpurun_out/pytest_stem.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --per-op --quick --no-cpu-baseline --no-e2e --sustained-seconds 3 > gpurun_out/bench_stem.json 2> gpurun_out/bench_stem.err
echo "bench rc=$?" >> gpurun_out/bench_stem.err
