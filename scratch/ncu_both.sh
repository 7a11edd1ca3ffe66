#!/bin/bash
# gpurun -- bash scratch/ncu_both.sh tag : plain run, ncu launch list of the same command, then one --set full capture of the level kernels
tag=${1:-run}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --headline-only --quick"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { tail -5 gpurun_out/${tag}_plain.err; exit 1; }
python scratch/bench_digest.py gpurun_out/${tag}_plain.json 2>/dev/null | head -3
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 5000 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/${tag}_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"mg_pre_kernel|mg_post_first_kernel|mg_post_second_kernel|mg_resid_restrict_kernel|spmv_grid_kernel|cg_update|rs_fd_rows_kernel" -c 24 -f -o gpurun_out/${tag}_full python bench.py --steps 1 --warmup 1 --no-cpu-baseline --headline-only --quick > gpurun_out/${tag}_full.log 2>&1
echo "full rc=$?"
