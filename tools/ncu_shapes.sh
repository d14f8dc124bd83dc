#!/bin/bash
# One ncu --set full capture of the scan kernel per bench shape at SF100 (run on the GPU box via gpurun):
#   tools/ncu_shapes.sh q1_style q5_topn ...   -> gpurun_out/r02_ncu_<shape>.ncu-rep + a CSV of the raw page
for shape in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:scan --launch-skip 6 --launch-count 1 -f -o gpurun_out/r02_ncu_$shape \
    python bench.py --headline-only --no-gate --shape $shape --steps 6 --warmup 3 > gpurun_out/r02_ncu_$shape.log 2>&1
  ncu -i gpurun_out/r02_ncu_$shape.ncu-rep --page raw --csv > gpurun_out/r02_ncu_${shape}_raw.csv 2>/dev/null
  tail -2 gpurun_out/r02_ncu_$shape.log | cut -c1-300
done
