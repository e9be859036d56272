set -x
timeout 600 python bench.py --steps 2 --warmup 1 --wide-steps 1 --wide-warmup 1 --no-cpu > gpurun_out/r02_bench_plain.log 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 2 --warmup 1 --wide-steps 1 --wide-warmup 1 --no-cpu --no-crosscheck > gpurun_out/r02_ncu_launches.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'pg_lowcard_agg|pg_wide_agg|wide_locate|wide_reorder_scatter|j_probe_dense|j_build_fast|wide_number_buckets|wide_gather_kernel|batch_gather' -c 20 -f -o gpurun_out/r02_full python bench.py --steps 1 --warmup 0 --wide-steps 1 --wide-warmup 0 --no-cpu --no-e2e --no-crosscheck > gpurun_out/r02_ncu_full.log 2>&1
timeout 300 python tools/profile_partition.py > gpurun_out/r02_part_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'part_' -c 3 -f -o gpurun_out/r02_part python tools/profile_partition.py > gpurun_out/r02_ncu_part.log 2>&1
ls -la gpurun_out/
