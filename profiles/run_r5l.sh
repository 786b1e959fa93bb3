#!/bin/bash
# lazy reference + block pipeline of `vtn normalize`: tests that touch them, then the CLI wall again
mkdir -p gpurun_out
python -m pytest tests/test_ref_lazy_gpu.py tests/test_cli_gpu.py tests/test_configs_gpu.py tests/test_vcf_gpu.py tests/test_writer_gpu.py -x -q -m gpu > gpurun_out/r5l_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5l_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 20 --warmup 5 --no-sharded > gpurun_out/r5l_bench.json 2> gpurun_out/r5l_bench.err
echo "rc=$?" >> gpurun_out/r5l_bench.err
nproc >> gpurun_out/r5l_bench.err
