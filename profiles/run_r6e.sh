#!/bin/bash
# compact-layout kernel with the staging window chosen per batch + full-wave generic launch: parity, then C2 / C4
mkdir -p gpurun_out
python -m pytest tests/test_bi_gpu.py tests/test_configs_gpu.py tests/test_fullsize_gpu.py tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/r6e_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6e_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for w in c2 c4; do
  python bench.py --workload $w --steps 20 --warmup 5 --resident-only > gpurun_out/r6e_$w.json 2> gpurun_out/r6e_$w.err
done
python bench.py --workload c4 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r6e_bench_c4.json 2> gpurun_out/r6e_bench_c4.err
