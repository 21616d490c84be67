#!/bin/bash
# round-2 GPU pass G: full GPU suite, smoke, default bench + reference arm (as the driver runs them)
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider 2>&1 | tail -150 ) > gpurun_out/pytest_g.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
( time python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err ) 2> gpurun_out/bench_g.time
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_g.json 2> gpurun_out/bench_ref_g.err
echo done > gpurun_out/done.txt
