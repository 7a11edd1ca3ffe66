#!/bin/bash
# gpurun --gpus N -- bash scratch/scale.sh "8 4 2" tag : bench at each N, digest
tag=${2:-scale}
for n in $1; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/${tag}_n$n.json 2> gpurun_out/${tag}_n$n.err
  python - <<PY
import json
txt=open("gpurun_out/${tag}_n$n.json").read()
ls=[l for l in txt.split("\n") if l.startswith("{")]
if not ls:
    print($n, "NO JSON"); print(open("gpurun_out/${tag}_n$n.err").read()[-1500:])
else:
    d=json.loads(ls[-1]); print($n, "ms", round(d["ms_per_step"],2), "krylov", d["config"]["krylov_iters_per_step"], "e2e", round(d["e2e"]["s_per_step"],4), d["parity"], "stdout lines", len(txt.strip().split("\n"))); print("   ", {k:(round(v["ms_per_step"],2), int(v["launches_per_step"])) for k,v in d["kernels"].items()})
PY
done
