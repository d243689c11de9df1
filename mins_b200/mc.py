"""Monte-Carlo sharding: independent simulation runs spread over the GPUs of one box.

The reference's Monte-Carlo driver is a sequential bash loop, one process per seed
(mins_eval/launch/scripts/sim_calibration.sh:18-47; CI matrix .github/workflows/evaluation.yml:62-104). Runs share no
state, so the partition needs NO data-path collective: run r goes to rank r mod world, every rank drives its own GPU
(one ml_handle per LiDAR id), and only the per-run result records (a few numbers: RMSE, NEES, timing) are gathered at the
end for the report, exactly like the evaluator merges `<seed>.txt` files. One process per GPU, launched by torchrun;
`torch.distributed` (gloo on CPU, nccl on GPUs) is plumbing for the gather and the barrier around timed regions.
"""
import os
import time

import numpy as np


def shard(n_runs, world, rank):
    """run indices owned by `rank`: r mod world == rank (round-robin keeps ranks within one run of each other)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_runs, world))


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def gather_records(local_records, world, rank):
    """all ranks receive every run's record, ordered by run index. Records are small picklable dicts with a 'run' key."""
    if world == 1:
        return sorted(local_records, key=lambda r: r["run"])
    import torch.distributed as dist
    bucket = [None] * world
    dist.all_gather_object(bucket, local_records)
    out = [r for part in bucket for r in part]
    return sorted(out, key=lambda r: r["run"])


def run_sharded(run_fn, n_runs, seeds=None, barrier=None):
    """Execute `run_fn(run_index, seed) -> dict` for this rank's shard; returns (all records, local wall seconds).

    `barrier` (callable or None) brackets the timed region; the caller takes the max over ranks of the returned time.
    """
    rank, _, world = dist_env()
    seeds = list(range(n_runs)) if seeds is None else list(seeds)
    if len(seeds) != n_runs:
        raise ValueError("one seed per run")
    mine = shard(n_runs, world, rank)
    if barrier:
        barrier()
    t0 = time.perf_counter()
    local = []
    for r in mine:
        rec = dict(run_fn(r, seeds[r]))
        rec["run"], rec["seed"], rec["rank"] = r, seeds[r], rank
        local.append(rec)
    if barrier:
        barrier()
    dt = time.perf_counter() - t0
    return gather_records(local, world, rank), dt


def summarize(records, keys):
    """mean +- std over runs of each key (the evaluator's table, mins_eval/src/run_comparison.cpp:92-180)."""
    out = {}
    for k in keys:
        v = np.array([r[k] for r in records if k in r and r[k] is not None], dtype=np.float64)
        out[k] = {"mean": float(v.mean()) if len(v) else None, "std": float(v.std()) if len(v) else None, "n": int(len(v))}
    return out
