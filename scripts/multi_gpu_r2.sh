#!/bin/bash
# Round-2 multi-GPU evidence on ONE box: bench.py at N = 1, 2, 4, 8 launched as the driver does, then the Monte-Carlo job of
# the host-C++ harness (64 seeds x 10 s of UD_Warehouse) at 1 / 2 / 4 / 8 GPUs with 4 processes per GPU. Outputs: gpurun_out/mg_*.
NG=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/mg_topo.txt 2>&1
nproc > gpurun_out/mg_nproc.txt
for N in 1 2 4 8; do
  [ $N -gt $NG ] && continue
  if [ $N = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 5 --no-aux --no-cpu-baseline > gpurun_out/mg_bench_g$N.json 2> gpurun_out/mg_bench_g$N.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py \
      --gpus $N --steps 20 --warmup 5 > gpurun_out/mg_bench_g$N.json 2> gpurun_out/mg_bench_g$N.err
  fi
  python scripts/bench_brief.py gpurun_out/mg_bench_g$N.json
done
for N in 1 2 4 8; do
  [ $N -gt $NG ] && continue
  python scripts/mc_sim_cpp.py --gpus $N --runs 64 --per-gpu 4 --duration 10 > gpurun_out/mg_mc64_g${N}x4.json 2> gpurun_out/mg_mc64_g${N}x4.err
  tail -c 700 gpurun_out/mg_mc64_g${N}x4.json; echo
done
