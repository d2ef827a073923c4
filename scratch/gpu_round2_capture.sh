#!/bin/bash
# What profiles/r02_* was made from, in one gpurun call on one B200 (every ncu run after the same command exited 0 plainly).
set -x
B="python bench.py"
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"
timeout 900 $B > gpurun_out/r02_bench_c2_1gpu.json 2> gpurun_out/r02_bench_c2_1gpu.err; echo "bench rc=$?"
LSG_JOIN_PATH=slab timeout 600 $B --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_c2_1gpu_slab.json 2> gpurun_out/r02_bench_c2_1gpu_slab.err; echo "slab rc=$?"
timeout 600 $B --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err; echo "ref rc=$?"
timeout 300 $B --workload c2g1 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-extra > gpurun_out/r02_bench_c2_one_group.json 2> gpurun_out/r02_bench_c2_one_group.err; echo "g1 rc=$?"
S="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extra"
timeout 200 $B $S > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 100 -c 120 --csv --log-file gpurun_out/r02_launches_c2.csv $B $S > gpurun_out/ncu1.log 2>&1; echo "ncu1 rc=$?"
LSG_JOIN_PATH=slab timeout 200 $B $S > gpurun_out/plain2.log 2>&1 && LSG_JOIN_PATH=slab ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 60 -c 80 --csv --log-file gpurun_out/r02_launches_c2_slab.csv $B $S > gpurun_out/ncu2.log 2>&1; echo "ncu2 rc=$?"
timeout 200 $B $S > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_join_tiles|k_count_keys|k_place_events" -s 9 -c 3 -o gpurun_out/r02_perm_full $B $S > gpurun_out/ncu3.log 2>&1; echo "ncu3 rc=$?"
LSG_JOIN_PATH=slab timeout 200 $B $S > gpurun_out/plain4.log 2>&1 && LSG_JOIN_PATH=slab ncu --set full --clock-control none --import-source on -k regex:"k_join_slab|k_bucket_scatter|k_bucket_count" -s 15 -c 5 -o gpurun_out/r02_slab_full $B $S > gpurun_out/ncu4.log 2>&1; echo "ncu4 rc=$?"
tail -3 gpurun_out/r02_pytest_gpu.log
