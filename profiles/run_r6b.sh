#!/bin/bash
# N=8: end-to-end scaling with and without the NUMA binding of each rank
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r6b_topo.txt 2>&1; nproc >> gpurun_out/r6b_topo.txt; lscpu | grep -i "numa\|socket\|model name" >> gpurun_out/r6b_topo.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for mode in bound unbound; do
  if [ $mode = unbound ]; then export VTN_BENCH_NO_NUMA=1; fi
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 5 --no-cpu-baseline --no-sharded --no-compare > gpurun_out/r6b_bench_n8_$mode.json 2> gpurun_out/r6b_bench_n8_$mode.err
  echo "rc=$?" >> gpurun_out/r6b_bench_n8_$mode.err
done
