CMD="python bench.py --steps 2 --warmup 3 --warmup-seconds 0 --no-cpu-baseline --no-aux"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_scan_tile -s 6 -c 3 -o gpurun_out/r2_tile_prof $CMD > gpurun_out/r2_tile_ncu.log 2>&1
tail -3 gpurun_out/r2_tile_ncu.log
