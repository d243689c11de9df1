"""Host side in the reference's language (C++): the UpdaterLidar/LidarHelper mirror over the C ABI
(mins_b200/host/*.hpp). CPU only: the queueing policy of UpdaterLidar.cpp:50-160 is driven with a scripted state and a
recording mock of the ABI; the link check proves the shim binds the real library's symbols."""
import os
import subprocess

from conftest import ROOT

HOST = os.path.join(ROOT, "mins_b200", "host")


def run(name, *args):
    return subprocess.run([os.path.join(HOST, name), *args], capture_output=True, text=True, check=True).stdout


def test_shim_links_against_the_real_abi():
    assert "ml_api bound: 23 entries" in run("shim_link_check")


def test_queueing_policy_trace():
    out = run("shim_policy_test")
    lines = [l for l in out.splitlines() if not l.startswith("#")]
    it = iter(lines)

    def expect(prefix):
        for l in it:
            if l.startswith(prefix):
                return l
        raise AssertionError(f"missing trace line starting with {prefix!r}\n{out}")

    expect("create h0")
    assert "raw=1 new=0 used=0" in expect("stacks")                  # no clones: wait (UpdaterLidar.cpp:70-71)
    # first usable scan initialises the map: ctor range filter -> downsample -> Build (UpdaterLidar.cpp:88-95)
    assert "t_rel=0.000" in expect("h0 scan_upload") and "range=[1.00,100.0]" in lines[2]
    expect("h0 scan_downsample leaf=0.30")
    expect("h0 map_build_from_scan")
    assert "raw=1 new=0 used=0" in expect("stacks")                  # 10.10 + 0.05 > newest clone: delayed (:78-82)
    assert "raw=0 new=2 used=0" in expect("stacks")
    # update(): exactly-at-clone scan has one pose block per time, calibration first in H_order (:181-214)
    u1 = expect("h0 scan_update")
    assert "hx=18 nI=1 nA=1 calib_col=0 cols=[6 | 12] at_clone=11" in u1 and "leaf=0.30" in u1 and "chi2=7.8147" in u1
    assert "rows=18 hx=18" in expect("EKFUpdate")
    u2 = expect("h0 scan_update")
    assert "hx=36 nI=4 nA=1" in u2 and "cols=[6 12 18 24 | 30] at_clone=01" in u2 and "tx=0.120" in u2
    expect("EKFUpdate")
    # all updates first, then registrations in order with the transform each update used (:118-124, LidarHelper.cpp:196-207)
    assert "tx=0.100" in expect("h0 scan_transform")
    expect("h0 map_add_scan ds=1 size=0.30")
    assert "tx=0.120" in expect("h0 scan_transform")
    expect("h0 map_add_scan ds=1 size=0.30")
    assert "raw=0 new=0 used=0" in expect("stacks")
    assert "n_acp=20 n_rej=2 last_up=10.12 anchor=10.00" in expect("chi")
    # re-anchor: old LiDAR frame -> newest clone's, decay options forwarded, last_up_time kept (LidarHelper.cpp:209-266)
    assert "tx=-0.300 now=0.300 decay=30.0/100.0 ds=1" in expect("h0 map_reanchor")
    assert "anchor=10.30 last_up=10.12" in expect("anchor")
    # soft "no rows": scan stays in the new stack (UpdaterLidar.cpp:398-399, :115-117); later dropped by the window
    expect("h0 scan_update")
    assert "raw=0 new=1 used=0" in expect("stacks")
    assert "raw=1 new=0 used=0" in expect("stacks")
    # stale map (> 1 s without registration): re-initialised from the next usable scan (:91)
    expect("h0 map_build_from_scan")
    assert "anchor=11.50 last_up=11.50 t_hist=6" in expect("anchor")
    assert out.count("map_reanchor") == 1
    # map_use_icp (LidarHelper.cpp:156-187): upload -> downsample -> ICP from the state's T_LtoM -> transform with the ICP
    # result -> point loop; the scan is then registered with the ICP transform (lidar->T_LtoM, LidarHelper.cpp:203)
    expect("h0 scan_upload")
    expect("h0 scan_downsample leaf=0.30")
    assert "tx=0.060 max_dist=50.0 seed=0" in expect("h0 scan_icp")
    assert "tx=0.560" in expect("h0 scan_transform")
    expect("h0 linsys hx=30")
    expect("EKFUpdate")
    expect("h0 map_add_scan")
    expect("icp_success=1")
    # ICP failed in update(): the point loop ran with the state's T_LtoM (tx = 0.120); register_scan retries transform_to_map
    # (upload -> downsample -> ICP again) and registers with the retry's result (LidarHelper.cpp:196-207)
    assert "tx=0.120" in expect("h0 scan_icp")
    assert "tx=0.120" in expect("h0 scan_transform")
    expect("EKFUpdate")
    expect("h0 scan_upload")
    assert "tx=0.120" in expect("h0 scan_icp")
    assert "tx=0.620" in expect("h0 scan_transform")
    expect("h0 map_add_scan")
    assert "icp_success=0 last_up=11.62" in expect("icp_success")
    # ICP fails twice: no map_add_scan, last_up_time stays (so the 1 s staleness re-initialisation can trigger, UpdaterLidar.cpp:91)
    tail = out[out.index("# ICP fails twice"):out.index("# raw_remove_motion_blur")]
    assert tail.count("scan_icp") == 2 and "map_add_scan" not in tail and "last_up=11.62" in tail and "raw=0 new=0 used=0" in tail
    # raw_remove_motion_blur (UpdaterLidar.cpp:82, LidarHelper.cpp:39-87): 5 knots over a 10 ms sweep at 1 m/s: the first knot
    # sits 1 cm behind the scan-time frame, the last one is the identity
    expect("h0 scan_upload")
    d = expect("h0 scan_deskew")
    assert "knots=5 total=1000 tx_first=-0.0100 tx_last=0.0000" in d
    expect("h0 scan_downsample")
    expect("h0 scan_transform")
    expect("h0 linsys")
    expect("h0 map_add_scan")
    # queued scans live in pooled pinned buffers, PointXYZ layout: few allocations for the ten scans fed, and the device is
    # handed stride-4 points
    assert int(expect("host_allocs=").split("=")[1]) <= 3 and "stride=3" not in out and "stride=4" in out


def test_kd_tree_facade_keeps_the_members_other_mins_code_uses():
    """Initializer.cpp:191-201,330-338 asks `ikd->tree->initialized()`, ROSPublisher.cpp:432-470 calls
    `ikd->tree->flatten(ikd->tree->Root_Node, pts, NOT_RECORD)`. mins_gpu::KD_TREE (host/lidar_gpu.hpp) offers those members over one
    ml_handle: traced here against the recording mock of the ABI."""
    out = run("shim_policy_test")
    tail = out[out.index("# KD_TREE facade"):]
    assert "initialized=0 root=0" in tail                                # after reset(): no map, Root_Node null
    assert "h0 map_build n=7" in tail and "initialized=1 root=1 size=7 validnum=7" in tail
    assert "h0 map_add n=3 downsample=1 size=0.25" in tail               # Add_Points(pts, true) with set_downsample_param(0.25)
    assert "live=10 flattened=10 x_last=9.0 intensity=0.25" in tail      # flatten() fills a PointVector of PointXYZI
