#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py tests/test_configs_gpu.py -x -q -m gpu -k "not filter_expression and not interval" > gpurun_out/r6v_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6v_tests.txt
N=3000000 python profiles/time_tools.py > gpurun_out/r6v_tools_3m.json 2> gpurun_out/r6v_tools.err
