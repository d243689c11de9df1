set -x
CMD="python bench.py --steps 2 --warmup 3 --warmup-seconds 0 --no-cpu-baseline --no-aux"
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err || exit 1
$CMD > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r1_launches_final.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_rows_reduce|k_knn|k_range_filter_bin|k_unit_alloc|k_query_scatter" -s 10 -c 10 -o gpurun_out/r1_final_prof $CMD > gpurun_out/ncu_full.log 2>&1
MI_IT=2 python scripts/map_insert_diag.py > /dev/null 2>&1 || exit 1
MI_IT=2 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1_map_launches.csv python scripts/map_insert_diag.py > /dev/null 2>&1
MI_IT=2 ncu --set full --clock-control none --import-source on -k regex:"k_map_merge|k_scan_lb|k_cell_count|k_cell_scatter|k_vox_|k_radix_pass|k_bbox_voxgrid|k_reanchor" -s 16 -c 24 -o gpurun_out/r1_map_prof python scripts/map_insert_diag.py > gpurun_out/ncu_map.log 2>&1
MB_WARM=3 MB_IT=2 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1_modeb_launches.csv python scripts/modeb_diag.py > /dev/null 2>&1
tail -c 600 gpurun_out/bench_final.json
