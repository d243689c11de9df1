# Round-2 final evidence capture for the one-kernel path (GPU box, one B200). Every ncu pass follows the same command having exited 0
# without ncu. Offline afterwards (no GPU): scripts/ncu_summary.py, scripts/ncu_lines.py on gpurun_out/r2_tile_prof.ncu-rep.
set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_final_pytest.log 2>&1; tail -3 gpurun_out/r2_final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_final_smoke.log 2>&1; tail -1 gpurun_out/r2_final_smoke.log
CMD="python bench.py --steps 2 --warmup 3 --warmup-seconds 0 --no-cpu-baseline --no-aux"
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err || exit 1
timeout 120 $CMD > /dev/null 2>&1 || exit 1
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_tile_launches.csv $CMD > /dev/null 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_scan_tile -s 6 -c 3 -o gpurun_out/r2_tile_prof $CMD > gpurun_out/r2_tile_ncu.log 2>&1
timeout 60 python scripts/timeline.py > gpurun_out/r2_tile_timeline_device.txt 2>&1
MINS_HOST_PROF=1 TL_HOST=1 timeout 60 python scripts/timeline.py > gpurun_out/r2_tile_timeline_host.txt 2>&1
timeout 120 python scripts/mc_sim_cpp.py --gpus 1 --runs 64 --per-gpu 8 --duration 10 > gpurun_out/r2_mc64_workers_g1x8.json 2> gpurun_out/r2_mc64_workers_g1x8.err
timeout 120 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err
python scripts/bench_brief.py gpurun_out/r2_bench_final.json
