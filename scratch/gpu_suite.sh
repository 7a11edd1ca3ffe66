#!/bin/bash
# Run on the GPU box (gpurun -- bash scratch/gpu_suite.sh [tag]): the whole -m gpu suite, then the bench, and a short digest.
tag=${1:-run}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider 2>&1 | tail -25 > gpurun_out/${tag}_tests.log
tail -6 gpurun_out/${tag}_tests.log
python bench.py --steps 5 --warmup 3 ${BENCH_FLAGS} > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
tail -5 gpurun_out/${tag}_bench.err
python scratch/bench_digest.py gpurun_out/${tag}_bench.json
