#!/bin/bash
# sliding-window lag_scan: parity, then C2 / C4 (both layouts)
mkdir -p gpurun_out
python -m pytest tests/test_normalize_gpu.py tests/test_bi_gpu.py tests/test_configs_gpu.py tests/test_fullsize_gpu.py tests/test_decompose_gpu.py -x -q -m gpu > gpurun_out/r6l_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6l_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for w in c2 c4; do
  for l in bi general; do
    python bench.py --workload $w --layout $l --steps 20 --warmup 5 --resident-only > gpurun_out/r6l_${w}_$l.json 2> gpurun_out/r6l_${w}_$l.err
  done
done
