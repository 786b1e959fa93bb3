#!/bin/bash
# round 2, first GPU call: parity of everything (incl. the compact biallelic layout), C2 bench on both layouts, L2 fetch
# granularity and register-budget variants of the bi kernel, launch list + one ncu --set full capture of the bi kernel.
mkdir -p gpurun_out
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
nvidia-smi -L > gpurun_out/r5a_gpu.txt
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r5a_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r5a_tests.log
python bench.py --steps 30 --warmup 5 > gpurun_out/r5a_bench_bi.json 2> gpurun_out/r5a_bench_bi.err
echo "bench rc=$?" >> gpurun_out/r5a_bench_bi.err
VTN_L2_FETCH=32 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/r5a_bench_bi_l2f32.json 2> gpurun_out/r5a_bench_bi_l2f32.err
VTN_L2_FETCH=128 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-compare > gpurun_out/r5a_bench_bi_l2f128.json 2> gpurun_out/r5a_bench_bi_l2f128.err
VTN_LIB_PATH=$PWD/vt_b200/exp/libvtn_mb7.so python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-compare > gpurun_out/r5a_bench_bi_mb7.json 2> gpurun_out/r5a_bench_bi_mb7.err
VTN_LIB_PATH=$PWD/vt_b200/exp/libvtn_mb7.so VTN_L2_FETCH=32 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-compare > gpurun_out/r5a_bench_bi_mb7_l2f32.json 2> gpurun_out/r5a_bench_bi_mb7_l2f32.err
CMD="python bench.py --steps 3 --warmup 3 --resident-only"
$CMD > gpurun_out/r5a_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r5a_launches.csv $CMD > gpurun_out/r5a_ncu_launches.log 2>&1
$CMD > gpurun_out/r5a_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:vtn_normalize_bi_kernel -s 4 -c 2 -o gpurun_out/r5a_prof $CMD > gpurun_out/r5a_ncu_full.log 2>&1
ls -la gpurun_out
