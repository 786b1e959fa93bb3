#!/bin/bash
mkdir -p gpurun_out
VTN_MULTI_CHUNK_REC=3000 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/r6h_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6h_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for slots in 2 3; do
  VTN_BENCH_C3S_SLOTS=$slots python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-compare > gpurun_out/r6h_n1_s$slots.json 2> gpurun_out/r6h_n1_s$slots.err
  VTN_BENCH_C3S_SLOTS=$slots python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-compare > gpurun_out/r6h_n2_s$slots.json 2> gpurun_out/r6h_n2_s$slots.err
done
VTN_BENCH_C3S_SLOTS=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-compare > gpurun_out/r6h_n2_s1.json 2> gpurun_out/r6h_n2_s1.err
