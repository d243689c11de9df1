#!/usr/bin/env python
"""BASELINE.json config 4: batched Monte-Carlo — N independent closed-loop IMU+LiDAR simulation runs sharded over the GPUs
of one box. One process per GPU (torchrun), run r on rank r mod world, no data-path collective; the per-run
records (ATE, NEES, counts) are gathered at the end like the reference's evaluator merges <seed>.txt files
(mins_eval/launch/scripts/sim_calibration.sh:18-47, .github/workflows/evaluation.yml:91-104).

  python scripts/mc_sim.py --runs 64 --duration 10
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 scripts/mc_sim.py --runs 64
  # several host processes per GPU (a run is host-bound: the EKF algebra is Python; the device is idle > 95 % of a run):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 16 ... scripts/mc_sim.py --runs 64 --gpus 2     # 8 ranks per GPU

Prints one JSON line on rank 0: aggregate scans/s and LiDAR points/s over the whole job (wall clock, max over ranks, bracketed
by barriers), the share of the time spent inside the LiDAR updater (host EKF algebra + device), and mean +- std of the ATE.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--runs", type=int, default=64)
    ap.add_argument("--duration", type=float, default=10.0)
    ap.add_argument("--icp", action="store_true", help="map_use_icp: true (the reference's simulation default)")
    ap.add_argument("--host-raycast", action="store_true", help="ray-cast the scans with numpy instead of ml_sim_raycast")
    ap.add_argument("--gpus", type=int, default=0, help="GPUs to use; ranks are spread round-robin (default: one per rank)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from mins_b200 import mc
    from mins_b200.lidar import OptionsLidar
    from mins_b200.sim.runner import SimConfig, run
    from mins_b200.sim.updater import GpuBackend

    rank, local_rank, world = mc.dist_env()
    if not torch.cuda.is_available():
        raise RuntimeError("mc_sim.py needs a CUDA device: the LiDAR path has no CPU fallback")
    n_gpus = args.gpus if args.gpus > 0 else min(world, torch.cuda.device_count())
    dev = local_rank % n_gpus
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("gloo")  # host-side plumbing only (barrier, gather of the records); ranks may share a GPU
    op = OptionsLidar(raw_do_downsample=True, map_use_icp=args.icp, map_icp_dist=50.0)

    def one(run_idx, seed):
        be = GpuBackend(max_scan_points=1 << 14, max_map_points=1 << 18, device=dev)
        try:
            cfg = SimConfig(seed=seed, duration=args.duration, op=op, scan_source="numpy" if args.host_raycast else "cuda")
            r = run(cfg, be)
            dev_ms = be.device_ms
        finally:
            be.close()
        keep = ("rmse_ori_deg", "rmse_pos_m", "nees_pos", "n_updates", "n_scans", "ldr_wall_s", "n_acc", "n_points", "icp_runs", "icp_ok")
        rec = {k: r[k] for k in keep}
        rec["device_ms"] = dev_ms
        return rec

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    one(-1, 10_000 + rank)  # warm-up run (library load, first allocations), not timed, not recorded
    recs, dt = mc.run_sharded(one, args.runs, barrier=barrier)
    t = torch.tensor([dt], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    wall = float(t[0])
    if rank == 0:
        scans = sum(r["n_scans"] for r in recs)
        pts = sum(r["n_points"] for r in recs)
        line = {"metric": "Monte-Carlo closed-loop runs: LiDAR scans/s over the whole job", "value": scans / wall, "unit": "scans/s",
                "points_per_s": pts / wall, "n_gpus": n_gpus, "ranks": world, "runs": args.runs, "sim_seconds_per_run": args.duration, "wall_s": wall,
                "scaling": "strong", "parallelism": f"{world} rank(s) on {n_gpus} GPU(s), run r on rank r mod {world}, no collective on the data path",
                "lidar_updater_share_of_wall": sum(r["ldr_wall_s"] for r in recs) / (wall * world),
                "device_share_of_wall": sum(r["device_ms"] for r in recs) * 1e-3 / (wall * world),
                "map_use_icp": bool(args.icp), "scan_source": "numpy" if args.host_raycast else "cuda (ml_sim_raycast)",
                "summary": mc.summarize(recs, ["rmse_ori_deg", "rmse_pos_m", "nees_pos", "n_updates"]),
                "per_run": [{k: (round(float(r[k]), 5) if isinstance(r[k], float) else r[k])
                             for k in ("seed", "rank", "rmse_ori_deg", "rmse_pos_m", "n_updates", "n_acc", "icp_runs", "icp_ok")} for r in recs],
                "host_cores": os.cpu_count()}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
