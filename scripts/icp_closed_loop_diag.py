"""GPU-box diagnostic: closed-loop accuracy with/without ICP pre-alignment, GPU backend vs CPU-oracle backend."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mins_b200.lidar import OptionsLidar
from mins_b200.sim.runner import SimConfig, run
from mins_b200.sim.updater import GpuBackend
from oracle.backend import OracleBackend

dur = float(sys.argv[1]) if len(sys.argv) > 1 else 6.0
for seed in (3, 4):
    for icp in (False, True):
        op = OptionsLidar(raw_do_downsample=True, map_use_icp=icp, map_icp_dist=50.0)
        g = GpuBackend(max_scan_points=1 << 14, max_map_points=1 << 18)
        t0 = time.time()
        r = run(SimConfig(seed=seed, duration=dur, op=op, scan_source="cuda"), g)
        g.close()
        print(f"GPU seed {seed} icp={icp}: rmse {r['rmse_pos_m']:.4f} m {r['rmse_ori_deg']:.3f} deg upd {r['n_updates']} acc {r['n_acc']}/{r['n_points']} "
              f"icp ok {r['icp_ok']}/{r['icp_runs']} iters {r['icp_iters']} map {r['map_points']} wall {time.time()-t0:.1f}s", flush=True)
op = OptionsLidar(raw_do_downsample=True, map_use_icp=True, map_icp_dist=50.0)
t0 = time.time()
r = run(SimConfig(seed=3, duration=min(dur, 4.0), op=op), OracleBackend())
print(f"CPU oracle seed 3 icp=True ({min(dur,4.0)} s): rmse {r['rmse_pos_m']:.4f} m {r['rmse_ori_deg']:.3f} deg icp ok {r['icp_ok']}/{r['icp_runs']} "
      f"iters {r['icp_iters']} wall {time.time()-t0:.1f}s", flush=True)
op = OptionsLidar(raw_do_downsample=True, map_use_icp=False)
r = run(SimConfig(seed=3, duration=min(dur, 4.0), op=op), OracleBackend())
print(f"CPU oracle seed 3 icp=False: rmse {r['rmse_pos_m']:.4f} m {r['rmse_ori_deg']:.3f} deg")
