#!/bin/bash
# smoke, ncu traffic of the current K1 (profiles/traffic.json), launch list, then the other configurations' lines
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r6c_smoke.log 2>&1
echo "rc=$?" >> gpurun_out/r6c_smoke.log
export VTN_BENCH_CACHE=/dev/shm/vtn_cache
python bench.py --steps 3 --warmup 3 --resident-only > gpurun_out/r6c_resident.json 2> gpurun_out/r6c_resident.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r6c_launches.csv python bench.py --steps 3 --warmup 3 --resident-only > gpurun_out/r6c_ncu1.log 2>&1
for w in c3 c4 c5; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r6c_bench_$w.json 2> gpurun_out/r6c_bench_$w.err
  echo "rc=$?" >> gpurun_out/r6c_bench_$w.err
done
