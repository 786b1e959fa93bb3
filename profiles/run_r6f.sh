#!/bin/bash
# pipelined chunks in vtn_multi_decompose_normalize: parity (2 GPUs) and the sharded C3 leg at N=1 (1 and 3 slots) and N=2
mkdir -p gpurun_out
python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/r6f_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6f_tests.txt
VTN_MULTI_CHUNK_REC=3000 python -m pytest tests/test_multi_gpu.py -x -q -m gpu >> gpurun_out/r6f_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6f_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for slots in 1 2 3; do
  VTN_BENCH_C3S_SLOTS=$slots python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-compare > gpurun_out/r6f_n1_s$slots.json 2> gpurun_out/r6f_n1_s$slots.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-compare > gpurun_out/r6f_n2_s3.json 2> gpurun_out/r6f_n2_s3.err
