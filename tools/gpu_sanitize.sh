#!/bin/bash
# compute-sanitizer over the round-2 kernels (seg, cell path, pack / stain, host pipe) on small test cases
mkdir -p gpurun_out
T="tests/test_cellpost.py tests/test_tilefuse.py tests/test_stain.py"
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest $T -m gpu -q -x -p no:cacheprovider -k "not finalize_tiled" > gpurun_out/sanitize_memcheck_cell.log 2>&1; echo "memcheck cell rc=$?" > gpurun_out/sanitize_rc.txt
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_seg.py -m gpu -q -x -p no:cacheprovider -k "not config2 and not full_mask and not otsu_mode" > gpurun_out/sanitize_memcheck_seg.log 2>&1; echo "memcheck seg rc=$?" >> gpurun_out/sanitize_rc.txt
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_image_cache.py -m gpu -q -x -p no:cacheprovider -k "fast_pack or extract" > gpurun_out/sanitize_memcheck_pack.log 2>&1; echo "memcheck pack rc=$?" >> gpurun_out/sanitize_rc.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_seg.py -m gpu -q -x -p no:cacheprovider -k "stage_closing or stage_median_other or cohort_batch or legacy" > gpurun_out/sanitize_racecheck_seg.log 2>&1; echo "racecheck seg rc=$?" >> gpurun_out/sanitize_rc.txt
for f in gpurun_out/sanitize_*.log; do tail -c 3000 $f > $f.tail; done
cat gpurun_out/sanitize_rc.txt
