#!/bin/bash
# staging window of the fast launch of the compact-layout kernel: 8 (default) / 16 / 24 bytes per record, on C2 and C4
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
for v in base bpr16 bpr24; do
  if [ $v = base ]; then unset VTN_LIB_PATH; else export VTN_LIB_PATH=$PWD/vt_b200/exp/libvtn_$v.so; fi
  for w in c2 c4; do
    python bench.py --workload $w --steps 20 --warmup 5 --resident-only > gpurun_out/r6d_${w}_$v.json 2> gpurun_out/r6d_${w}_$v.err
  done
done
