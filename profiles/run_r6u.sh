#!/bin/bash
# context creation overlapped with the host's FASTA / header / first-contig work; next contig prefetched
mkdir -p gpurun_out
python -m pytest tests/test_ref_lazy_gpu.py tests/test_cli_gpu.py -x -q -m gpu -k "normalize or c1 or lazy or host_half" > gpurun_out/r6u_tests.txt 2>&1
echo "rc=$?" >> gpurun_out/r6u_tests.txt
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 5 --warmup 3 --no-sharded --no-compare > gpurun_out/r6u_bench.json 2> gpurun_out/r6u_bench.err
echo "rc=$?" >> gpurun_out/r6u_bench.err
