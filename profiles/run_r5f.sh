#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests/test_configs_gpu.py -x -q ) > gpurun_out/r5f_config_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r5f_config_tests.log
python bench.py --workload c5 --steps 5 --warmup 3 > gpurun_out/r5f_bench_c5.json 2> gpurun_out/r5f_bench_c5.err
echo "rc=$?" >> gpurun_out/r5f_bench_c5.err
