# Round-2 evidence capture (GPU box, one B200). Every ncu pass follows the same command having exited 0 without ncu.
set -x
CMD="python bench.py --steps 2 --warmup 3 --warmup-seconds 0 --no-cpu-baseline --no-aux"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err || exit 1
timeout 300 $CMD > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > /dev/null 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_rows_reduce|k_front_knn" -s 8 -c 6 -o gpurun_out/r2_prof $CMD > gpurun_out/r2_ncu_full.log 2>&1
python scripts/ncu_summary.py gpurun_out/r2_prof.ncu-rep gpurun_out/r2_ncu_summary.json > gpurun_out/r2_ncu_summary.txt 2>&1
python scripts/ncu_lines.py gpurun_out/r2_prof.ncu-rep k_rows_reduce > gpurun_out/r2_lines_rows.txt 2>&1
python scripts/ncu_lines.py gpurun_out/r2_prof.ncu-rep k_front_knn > gpurun_out/r2_lines_front_knn.txt 2>&1
timeout 120 python scripts/timeline.py > gpurun_out/r2_timeline_device.txt 2>&1
TL_HOST=1 timeout 120 python scripts/timeline.py > gpurun_out/r2_timeline_host.txt 2>&1
cuobjdump -sass mins_b200/libmins_lidar.so | grep -E "Function : .*k_front_knnILi5ELb1|UBLKCP|SYNCS" | head -40 > gpurun_out/r2_sass_tma.txt
tail -c 400 gpurun_out/r2_bench_final.json
