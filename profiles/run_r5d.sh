#!/bin/bash
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r5d_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r5d_tests.log
python bench.py --steps 30 --warmup 5 > gpurun_out/r5d_bench_bi.json 2> gpurun_out/r5d_bench_bi.err
echo "bench rc=$?" >> gpurun_out/r5d_bench_bi.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r5d_bench_ref.json 2> gpurun_out/r5d_bench_ref.err
