#!/bin/bash
# A/B of the flat kernel's shared-memory plan knobs (slab stages vs staging buffers vs tiles per slab)
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 600 python bench.py --model ${MODEL:-resnet34} --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/knob_$tag.json 2> gpurun_out/knob_$tag.err
  python -c "import json;d=json.load(open('gpurun_out/knob_$tag.json'));print('$tag',d['value'],d['ms_per_step'],d['roofline']['achieved'])"
  grep "^op" gpurun_out/knob_$tag.err | head -20 | awk '{printf "%s:%s ", $2,$5}'; echo
}
run base X=1
run nout1 WSI_FLAT_NOUT=1
run t1 WSI_FLAT_T=1
run t1nout1 WSI_FLAT_T=1 WSI_FLAT_NOUT=1
MODEL=inceptionv4 run iv4_base X=1
MODEL=inceptionv4 run iv4_nout1 WSI_FLAT_NOUT=1
