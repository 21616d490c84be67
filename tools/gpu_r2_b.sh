#!/bin/bash
# round-2 GPU pass B: cell-path finalize tests + timing, full GPU suite again (parity bounds), CLI tests
mkdir -p gpurun_out
( time python -m pytest tests/test_cellpost.py tests/test_tilefuse.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -40 ) > gpurun_out/pytest_cellpost.log 2>&1
timeout 300 python tools/cellpost_bench.py > gpurun_out/cellpost_bench.json 2> gpurun_out/cellpost_bench.err
( time python -m pytest tests -m gpu -q -p no:cacheprovider --deselect tests/test_cellpost.py --deselect tests/test_tilefuse.py 2>&1 | tail -30 ) > gpurun_out/pytest_gpu.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo done > gpurun_out/done.txt
