#!/bin/bash
# compressed I/O (BGZF) in the tools
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -m gpu -k "bcf or bgzf or c1 or output_name" > gpurun_out/r5q_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5q_tests.txt
