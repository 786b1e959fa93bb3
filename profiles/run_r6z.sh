#!/bin/bash
# end-of-round verification on 2 GPUs: whole GPU suite, smoke, N=1 and N=2 bench lines, reference arm
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r6z_gpus.txt; nproc >> gpurun_out/r6z_gpus.txt; numactl -H >> gpurun_out/r6z_gpus.txt 2>&1; nvidia-smi topo -m >> gpurun_out/r6z_gpus.txt 2>&1
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r6z_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r6z_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r6z_smoke.log 2>&1
echo "rc=$?" >> gpurun_out/r6z_smoke.log
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 20 --warmup 5 > gpurun_out/r6z_bench_n1.json 2> gpurun_out/r6z_bench_n1.err
echo "rc=$?" >> gpurun_out/r6z_bench_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r6z_bench_n2.json 2> gpurun_out/r6z_bench_n2.err
echo "rc=$?" >> gpurun_out/r6z_bench_n2.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r6z_bench_ref.json 2> gpurun_out/r6z_bench_ref.err
echo "rc=$?" >> gpurun_out/r6z_bench_ref.err
