#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -q -m gpu -k "stress_file" > gpurun_out/r7c_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r7c_tests.txt
