#!/bin/bash
# compressed I/O (BGZF) in the tools
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -m gpu -k "not filter_expression" > gpurun_out/r5p_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5p_tests.txt
