#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py tests/test_configs_gpu.py tests/test_annot_gpu.py -x -q -m gpu  > gpurun_out/r6q_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6q_tests.txt
python profiles/time_tools.py > gpurun_out/r6q_tools.json 2> gpurun_out/r6q_tools.err
N=3000000 python profiles/time_tools.py > gpurun_out/r6q_tools_3m.json 2>> gpurun_out/r6q_tools.err
