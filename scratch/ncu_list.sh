#!/bin/bash
# gpurun -- bash scratch/ncu_list.sh tag : plain run, then the ncu launch list of the same command
tag=${1:-run}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --headline-only --quick"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { tail -5 gpurun_out/${tag}_plain.err; exit 1; }
python scratch/bench_digest.py gpurun_out/${tag}_plain.json | head -3
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/${tag}_launches.csv
