#!/bin/bash
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r5h_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r5h_tests.log
VTN_BENCH_SEED_RANK=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sharded --resident-only > gpurun_out/r5h_bench_seed1.json 2> gpurun_out/r5h_bench_seed1.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sharded --resident-only > gpurun_out/r5h_bench_seed0.json 2> gpurun_out/r5h_bench_seed0.err
