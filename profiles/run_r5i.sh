#!/bin/bash
mkdir -p gpurun_out
( python -m pytest tests/test_zz_no_alloc.py -x -q ) > gpurun_out/r5i_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/r5i_noalloc_alone.log
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r5i_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r5i_tests.log
