#!/bin/bash
# final verification of the round's last state on one GPU: whole GPU suite, smoke, default bench, reference arm
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r7a_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r7a_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r7a_smoke.log 2>&1
echo "rc=$?" >> gpurun_out/r7a_smoke.log
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py > gpurun_out/r7a_bench_n1.json 2> gpurun_out/r7a_bench_n1.err
echo "rc=$?" >> gpurun_out/r7a_bench_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r7a_bench_ref.json 2> gpurun_out/r7a_bench_ref.err
echo "rc=$?" >> gpurun_out/r7a_bench_ref.err
