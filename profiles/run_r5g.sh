#!/bin/bash
# 2-GPU box: the whole GPU suite (incl. the multi-device tests on both GPUs), bench at N=1 and N=2 with the sharded C3 leg
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
nvidia-smi -L > gpurun_out/r5g_gpus.txt; nvidia-smi topo -m >> gpurun_out/r5g_gpus.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r5g_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r5g_tests.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r5g_bench_n1.json 2> gpurun_out/r5g_bench_n1.err
echo "rc=$?" >> gpurun_out/r5g_bench_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r5g_bench_n2.json 2> gpurun_out/r5g_bench_n2.err
echo "rc=$?" >> gpurun_out/r5g_bench_n2.err
