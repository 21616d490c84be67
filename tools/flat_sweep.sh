#!/bin/bash
# sweep the flat-conv shared-memory plan (bench only): T, nOut, weight residency
mkdir -p gpurun_out
for cfg in "2 0 -1" "1 0 -1" "1 1 -1" "2 1 -1" "1 2 0" "2 2 0" "1 1 0"; do
  set -- $cfg
  export WSI_FLAT_T=$1
  if [ "$2" != "0" ]; then export WSI_FLAT_NOUT=$2; else unset WSI_FLAT_NOUT; fi
  if [ "$3" != "-1" ]; then export WSI_FLAT_RESIDENT=$3; else unset WSI_FLAT_RESIDENT; fi
  timeout 200 python bench.py --steps 20 --warmup 3 --per-op --no-cpu-baseline --no-e2e > gpurun_out/sweep.json 2> gpurun_out/sweep.err
  echo "T=$1 nOut=$2 resident=$3 rc=$? $(python -c "import json; d=json.load(open('gpurun_out/sweep.json')); print(round(d['value']), round(d['ms_per_step'],3))")"
  grep "^op" gpurun_out/sweep.err | sed -n '3,4p;11,12p' | awk '{print "   ", $1,$2,$3,$5}'
done
