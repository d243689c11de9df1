timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_full.py -m gpu -x -q 2>&1 | tail -6
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-aux > gpurun_out/r2_bench_t6.json 2> gpurun_out/r2_bench_t6.err
python scripts/bench_brief.py gpurun_out/r2_bench_t6.json
python -c "
import json
d=json.loads(open('gpurun_out/r2_bench_t6.json').read().strip().splitlines()[-1]); print(d['e2e']['by_input'], d['e2e']['step_ms_min_median_max'])"
timeout 120 python scripts/timeline.py 2>&1 | head -8
MINS_HOST_PROF=1 TL_HOST=1 timeout 120 python scripts/timeline.py 2>&1 | grep -E 'host prof|scan 2' -A7 | head -9
