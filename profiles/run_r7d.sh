#!/bin/bash
# after the last host-side changes (BCF encoder fix, self-test hooks): the command-line tests once more
mkdir -p gpurun_out
python -m pytest tests/test_cli_gpu.py tests/test_ref_lazy_gpu.py tests/test_variant_manip_gpu.py -x -q -m gpu > gpurun_out/r7d_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r7d_tests.txt
