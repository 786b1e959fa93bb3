#!/bin/bash
# -f filter subset
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py tests/test_annot_gpu.py -x -q -m gpu > gpurun_out/r5o_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5o_tests.txt
