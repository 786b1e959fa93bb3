#!/bin/bash
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 20 --warmup 5 --no-sharded > gpurun_out/r5j_bench.json 2> gpurun_out/r5j_bench.err
echo "rc=$?" >> gpurun_out/r5j_bench.err
