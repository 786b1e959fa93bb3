#!/bin/bash
# ncu --set full capture of the dominant kernel (compact-layout K1) for roofline.traffic
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 3 --warmup 3 --resident-only > gpurun_out/r6i_resident.json 2> gpurun_out/r6i_resident.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:vtn_normalize_bi_kernel -s 4 -c 4 -f -o gpurun_out/r6i_k1 python bench.py --steps 3 --warmup 3 --resident-only > gpurun_out/r6i_ncu.log 2>&1
echo "rc=$?" >> gpurun_out/r6i_ncu.log
