#!/bin/bash
# verbatim pass-through in put_tail for every tool: CLI parity, then the tools' wall against the reference
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py tests/test_configs_gpu.py tests/test_annot_gpu.py -x -q -m gpu > gpurun_out/r6n_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6n_tests.txt
python profiles/time_tools.py > gpurun_out/r6n_tools.json 2> gpurun_out/r6n_tools.err
