#!/bin/bash
# What profiles/r02_*_v3 / r02_ncu_* was made from at the end of round 2, in one gpurun call on one B200 (every ncu run
# after the same command exited 0 plainly).
set -x
B="python bench.py"
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu_v3.log 2>&1; echo "pytest rc=$?"
timeout 900 $B > gpurun_out/r02_bench_c2_1gpu_v3.json 2> gpurun_out/r02_bench_c2_1gpu_v3.err; echo "bench rc=$?"
timeout 600 $B --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_reference_arm_v3.json 2> gpurun_out/r02_bench_reference_arm_v3.err; echo "ref rc=$?"
S="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-extra"
timeout 200 $B $S > gpurun_out/plain.log 2>&1 && ncu --metrics $M --clock-control none -s 100 -c 120 --csv --log-file gpurun_out/r02_launches_c2_v3.csv $B $S > gpurun_out/ncu1.log 2>&1; echo "ncu1 rc=$?"
timeout 200 $B $S > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_join_tiles|k_count_keys|k_place_events" -s 9 -c 3 -f -o gpurun_out/r02_perm_full_v3 $B $S > gpurun_out/ncu3.log 2>&1; echo "ncu3 rc=$?"
S3="--steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-extra --workload c3"
timeout 300 $B $S3 > gpurun_out/plain5.log 2>&1 && ncu --metrics $M --clock-control none -s 150 -c 60 --csv --log-file gpurun_out/r02_launches_c3_v3.csv $B $S3 > gpurun_out/ncu5.log 2>&1; echo "ncu5 rc=$?"
timeout 300 $B $S3 > gpurun_out/plain6.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_join_ring|k_radix_scatter|k_radix_hist|k_event_keys" -s 24 -c 8 -f -o gpurun_out/r02_c3_full_v3 $B $S3 > gpurun_out/ncu6.log 2>&1; echo "ncu6 rc=$?"
tail -3 gpurun_out/r02_pytest_gpu_v3.log
ls -la gpurun_out/*.ncu-rep
