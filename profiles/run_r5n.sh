#!/bin/bash
# decompose_blocksub -a on the device + CLI
mkdir -p gpurun_out
python -m pytest tests/test_blocksub_gpu.py tests/test_cli_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r5n_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5n_tests.txt
