#!/bin/bash
mkdir -p gpurun_out
export ASAN_OPTIONS=detect_leaks=0:protect_shadow_gap=0:abort_on_error=0:halt_on_error=1:log_path=gpurun_out/r5c_asan_report
LD_PRELOAD=/usr/lib/gcc/x86_64-linux-gnu/13/libasan.so VTN_LIB_PATH=$PWD/vt_b200/exp/libvtn_asan.so timeout 900 python -m pytest tests/test_bi_gpu.py tests/test_blocksub_gpu.py -x -q -s > gpurun_out/r5c_asan.log 2>&1; echo "rc=$?" >> gpurun_out/r5c_asan.log
