#!/bin/bash
# compressed I/O (BGZF) in the tools
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py -x -q -m gpu -k "uniq or blocksub_cli_small or annot" > gpurun_out/r5s_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r5s_tests.txt
