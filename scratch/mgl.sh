#!/bin/bash
# gpurun --gpus 2 -- bash scratch/mgl.sh : global-level count / push grid at 2 GPUs
for cfg in "RS_MG_GLOBAL_LEVELS=2" "RS_MG_GLOBAL_LEVELS=4" "RS_MG_PUSH_BLOCKS=32" "RS_MG_PUSH_BLOCKS=8"; do
  env $cfg python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --headline-only --quick > gpurun_out/mgl.json 2> gpurun_out/mgl.err
  echo "$cfg"; python scratch/bench_digest.py gpurun_out/mgl.json 2>/dev/null | grep -E "ms_per_step|halo|mg_coarse"
done
