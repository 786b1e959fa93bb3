#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_blocksub_gpu.py -x -q > gpurun_out/r5b_blocksub_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r5b_blocksub_alone.log
MALLOC_CHECK_=3 python -m pytest tests/test_bi_gpu.py tests/test_blocksub_gpu.py -x -q > gpurun_out/r5b_mcheck.log 2>&1; echo "rc=$?" >> gpurun_out/r5b_mcheck.log
timeout 600 cuda-gdb -batch -ex "set pagination off" -ex "set cuda break_on_launch none" -ex run -ex bt -ex "info threads" --args python -m pytest tests/test_bi_gpu.py tests/test_blocksub_gpu.py -x -q > gpurun_out/r5b_gdb.log 2>&1
python -m pytest tests -m gpu -q --deselect tests/test_blocksub_gpu.py > gpurun_out/r5b_rest.log 2>&1; echo "rc=$?" >> gpurun_out/r5b_rest.log
