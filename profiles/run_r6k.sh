#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -m gpu -k "in_process or stdin_to_stdout" > gpurun_out/r6k_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6k_tests.txt
