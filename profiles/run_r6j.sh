#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -m gpu -k "block_reader or window_zero" > gpurun_out/r6j_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6j_tests.txt
