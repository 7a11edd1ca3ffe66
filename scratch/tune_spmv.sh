#!/bin/bash
# compile SpMV variants on the GPU box and time the 256^3 bench with each
for cfg in "256 4 2304" "256 3 2304" "128 4 1152" "128 6 1152" "256 6 2304" "512 2 4608" "128 8 1152"; do
  set -- $cfg
  RS_NVCC_DEFS="-DRS_TMA_ROWS=$1 -DRS_TMA_STAGES=$2 -DRS_TMA_CAP=$3" python -m rootsolve_b200.build --force > /dev/null 2>&1 || { echo "build failed $cfg"; continue; }
  python bench.py --steps 2 --warmup 2 --no-cpu-baseline --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', round(d['ms_per_step'],1), round(d['roofline']['achieved']), round(d['roofline']['avg_launch_ms'],4))"
done
