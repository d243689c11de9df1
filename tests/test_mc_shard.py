"""Monte-Carlo sharding on CPU: the partition is exact and disjoint, and a world_size-2 gloo job gathers
every run's record in run order on every rank. No GPU, no data-path collective."""
import os
import subprocess
import sys
import textwrap

import pytest

from conftest import ROOT


@pytest.mark.parametrize("n_runs,world", [(64, 1), (64, 2), (64, 4), (64, 8), (7, 4), (3, 8), (0, 2)])
def test_shard_is_a_partition(n_runs, world):
    from mins_b200.mc import shard
    parts = [shard(n_runs, world, r) for r in range(world)]
    flat = sorted(x for p in parts for x in p)
    assert flat == list(range(n_runs))
    assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        shard(4, 2, 2)


def test_summarize():
    from mins_b200.mc import summarize
    s = summarize([{"run": 0, "rmse": 1.0}, {"run": 1, "rmse": 3.0}, {"run": 2}], ["rmse"])
    assert s["rmse"] == {"mean": 2.0, "std": 1.0, "n": 2}


def test_two_rank_gloo_job_gathers_all_runs(tmp_path):
    script = tmp_path / "job.py"
    script.write_text(textwrap.dedent(f"""
        import sys, json, os
        sys.path.insert(0, {ROOT!r})
        import torch.distributed as dist
        from mins_b200 import mc
        dist.init_process_group("gloo")
        rank, _, world = mc.dist_env()
        def run(i, seed):
            return {{"value": seed * seed, "pid": os.getpid()}}
        recs, dt = mc.run_sharded(run, 9, seeds=[10 + i for i in range(9)], barrier=dist.barrier)
        assert [r["run"] for r in recs] == list(range(9))
        assert [r["value"] for r in recs] == [(10 + i) ** 2 for i in range(9)]
        assert [r["rank"] for r in recs] == [i % world for i in range(9)]
        assert len({{r["pid"] for r in recs}}) == world
        if rank == 0:
            print("OK", json.dumps(mc.summarize(recs, ["value"])))
        dist.destroy_process_group()
    """))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29611", str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "OK " in out.stdout
