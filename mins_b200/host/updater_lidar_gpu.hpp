// UpdaterLidar over the C ABI — mirrors mins/src/update/lidar/UpdaterLidar.{h,cpp}. Included by lidar_gpu.hpp.
#pragma once

namespace mins_gpu {

class UpdaterLidar {
public:
  // UpdaterLidar::UpdaterLidar(shared_ptr<State>) (UpdaterLidar.cpp:42-48): one map, one statistics block and one time
  // history per LiDAR id. Capacities size the device buffers once (nothing is allocated on the per-scan path).
  UpdaterLidar(StateView *state, const OptionsLidar &op, const ml_api &api, int max_scan_points = 1 << 18, int max_map_points = 1 << 21,
               int device = 0)
      : op(op), state(state), api(api), cap_scan(max_scan_points) {
    for (int i = 0; i < op.max_n; i++) {
      Chi.emplace_back();
      t_hist[i] = std::deque<double>();
      auto ikd = std::make_shared<iKDDATA>();
      ikd->id = i;
      ml_config cfg;
      cfg.device = device, cfg.max_scan_points = max_scan_points, cfg.max_map_points = max_map_points;
      cfg.max_grid_cells = 0, cfg.grid_cell = 0.f;
      ml_handle *h = nullptr;
      if (api.create(&cfg, &h) != ML_OK)
        die("ml_create", nullptr);
      ikd->tree = std::make_shared<KD_TREE>(api, h);
      ikd->tree->set_downsample_param((float)op.map_downsample_size);
      ikd_data.push_back(ikd);
    }
  }
  ~UpdaterLidar() {
    stack_lidar_raw.clear(), stack_lidar_new.clear(), stack_lidar_used.clear(); // scans return their buffers to the pool
    for (float *b : pool)
      api.host_free(b);
    for (auto &ikd : ikd_data)
      api.destroy(ikd->tree->h);
  }
  UpdaterLidar(const UpdaterLidar &) = delete;
  UpdaterLidar &operator=(const UpdaterLidar &) = delete;

  // LiDARData::LiDARData's storage: a pooled pinned buffer of max_scan_points PointXYZ; the incoming points (n x stride floats)
  // are copied once, here, off the update's critical path. The range filter of the constructor runs on the device later.
  std::shared_ptr<LiDARData> make_scan(int id, double stamp, const float *xyz, int n, int stride) {
    if (n > cap_scan)
      die("feed_measurement: more points than max_scan_points", nullptr);
    float *buf = nullptr;
    if (!pool.empty()) {
      buf = pool.back();
      pool.pop_back();
    } else {
      buf = (float *)api.host_alloc(sizeof(float) * 4 * (size_t)cap_scan);
      if (!buf)
        die("ml_host_alloc", nullptr);
    }
    for (int i = 0; i < n; i++) {
      buf[4 * i + 0] = xyz[(size_t)stride * i + 0], buf[4 * i + 1] = xyz[(size_t)stride * i + 1], buf[4 * i + 2] = xyz[(size_t)stride * i + 2];
      buf[4 * i + 3] = 1.f;
    }
    std::shared_ptr<LiDARData> scan(new LiDARData(), [this](LiDARData *d) {
      if (d->xyz)
        pool.push_back(d->xyz);
      delete d;
    });
    scan->time = stamp, scan->id = id, scan->xyz = buf, scan->n = n;
    return scan;
  }

  // UpdaterLidar::feed_measurement (UpdaterLidar.cpp:50-104). `stamp` in seconds, xyz = n x stride floats (sensor frame).
  void feed_measurement(int id, double stamp, const float *xyz, int n, int stride = 3) {
    if (FT < 0)
      FT = stamp; // first ever measurement time = reference time of the intensity stamps (LidarTypes.h)
    auto &th = t_hist.at(id);
    if (th.size() > 100)
      th.pop_front();
    th.push_back(stamp);
    stack_lidar_raw.push_back(make_scan(id, stamp, xyz, n, stride));

    for (auto it = stack_lidar_raw.begin(); it != stack_lidar_raw.end();) {
      if (state->clones_empty())
        break;
      const double t = (*it)->time + state->lidar_dt((*it)->id);
      if (t < state->oldest_clone_time()) { // older than the window: drop
        it = stack_lidar_raw.erase(it);
        continue;
      }
      if (t + 0.05 > state->newest_clone_time()) { // newer than the newest clone (+0.05 s delay): wait
        ++it;
        continue;
      }
      auto &ikd = ikd_data.at((*it)->id);
      // map not initialised, or no successful registration for more than 1 s: (re)build it from this scan
      if (!api.map_initialized(ikd->tree->h) || ikd->last_up_time + 1 < (*it)->time) {
        if (!init_map_local(*it, ikd)) { // de-skew poses not available yet: keep it (UpdaterLidar.cpp:82-85)
          ++it;
          continue;
        }
        it = stack_lidar_raw.erase(it);
        continue;
      }
      stack_lidar_new.push_back(*it);
      it = stack_lidar_raw.erase(it);
    }
    erase_invalid_dataset();
  }

  // Initializer::init_lidar_sim (mins/src/init/Initializer.cpp:308-340; the Initializer is a friend of the reference's
  // UpdaterLidar): the dense ground-truth scan becomes the map as it is - LiDARData ctor (range filter) + init_map_local, no
  // down-sampling - and sets the reference time FT.
  void init_map_gt(int id, double stamp, const float *xyz, int n, int stride = 3) {
    if (FT < 0)
      FT = stamp;
    auto scan = make_scan(id, stamp, xyz, n, stride);
    auto &ikd = ikd_data.at(id);
    ck(api.scan_upload(ikd->tree->h, scan->xyz, 4, scan->size(), (float)(scan->time - FT), (float)op.min_range, (float)op.max_range,
                       nullptr),
       "ml_scan_upload", ikd->tree->h);
    resident[ikd->id] = nullptr;
    ck(api.map_build_from_scan(ikd->tree->h), "ml_map_build_from_scan", ikd->tree->h);
    ikd->time = stamp;
    ikd->last_up_time = stamp;
  }

  // UpdaterLidar::try_update (UpdaterLidar.cpp:106-125)
  void try_update() {
    for (auto it = stack_lidar_new.begin(); it != stack_lidar_new.end();) {
      if (update(*it, ikd_data.at((*it)->id), Chi.at((*it)->id))) {
        stack_lidar_used.push_back(*it);
        it = stack_lidar_new.erase(it);
      } else {
        ++it; // currently unusable: keep it
      }
    }
    for (auto it = stack_lidar_used.begin(); it != stack_lidar_used.end();) {
      register_scan(*it, ikd_data.at((*it)->id));
      it = stack_lidar_used.erase(it);
    }
  }

  // UpdaterLidar::propagate_map_frame (UpdaterLidar.cpp:147-160)
  void propagate_map_frame() {
    for (auto &ikd : ikd_data) {
      if (!api.map_initialized(ikd->tree->h))
        continue;
      if ((int)state->n_clones() < state->intr_order() + 1)
        continue;
      if (ikd->time + state->lidar_dt(ikd->id) <= state->oldest_3rd_clone_time())
        propagate_map_to_newest_clone(ikd);
    }
  }

  std::vector<UpdaterStatistics> Chi;             // UpdaterLidar.h: Chi
  std::map<int, std::deque<double>> t_hist;       // UpdaterLidar.h: t_hist
  std::vector<std::shared_ptr<iKDDATA>> ikd_data; // UpdaterLidar.h: ikd_data
  double FT = -1;
  std::vector<std::shared_ptr<LiDARData>> stack_lidar_raw, stack_lidar_new, stack_lidar_used; // friend Initializer touches these
  ml_counters last_counters{};
  int n_updates = 0; // successful EKF updates so far
  bool last_icp_success = true; // LiDARData::icp_success of the last update (LidarTypes.h)
  OptionsLidar op;              // the reference reads state->op->lidar; here the updater holds its copy

private:
  struct Pending { // what register_scan needs from update(): LiDARData::T_LtoM and LiDARData::icp_success (LidarTypes.h)
    float T_LtoM[16];
    bool icp_success = true;
  };
  std::map<const LiDARData *, Pending> pending;
  std::map<int, const LiDARData *> resident; // which scan currently sits in each handle's device buffers

  void die(const char *what, ml_handle *h) {
    std::fprintf(stderr, "\033[31m[LiDAR-GPU] %s failed: %s\033[0m\n", what, h ? api.last_error(h) : "(no handle)");
    std::exit(EXIT_FAILURE); // the reference's behaviour for unrecoverable errors (UpdaterLidar.cpp:210)
  }
  void ck(int rc, const char *what, ml_handle *h) {
    if (rc < 0)
      die(what, h);
  }

  // LidarHelper::init_map_local (LidarHelper.cpp:103-113) on a raw scan: constructor's range filter, optional
  // down-sampling (UpdaterLidar.cpp:88), then Build
  bool init_map_local(const std::shared_ptr<LiDARData> &scan, std::shared_ptr<iKDDATA> &ikd) {
    if (!upload(scan, ikd, /*downsample=*/op.raw_do_downsample))
      return false;
    ck(api.map_build_from_scan(ikd->tree->h), "ml_map_build_from_scan", ikd->tree->h);
    ikd->time = scan->time;
    ikd->last_up_time = scan->time;
    return true;
  }

  // LiDARData ctor + [remove_motion_blur] + [downsample] on the device (UpdaterLidar.cpp:61,82,88). Returns false when the
  // de-skew poses are not available yet (the reference keeps the scan and retries, UpdaterLidar.cpp:82-85).
  bool upload(const std::shared_ptr<LiDARData> &scan, std::shared_ptr<iKDDATA> &ikd, bool downsample) {
    std::vector<float> knots;
    if (op.raw_remove_motion_blur && !deskew_knots(scan, knots))
      return false;
    ck(api.scan_upload(ikd->tree->h, scan->xyz, 4, scan->size(), (float)(scan->time - FT), (float)op.min_range, (float)op.max_range,
                       nullptr),
       "ml_scan_upload", ikd->tree->h);
    if (!knots.empty())
      ck(api.scan_deskew(ikd->tree->h, knots.data(), (int)(knots.size() / 12), sweep_points(scan)), "ml_scan_deskew", ikd->tree->h);
    if (downsample)
      ck(api.scan_downsample(ikd->tree->h, (float)op.raw_downsample_size, nullptr), "ml_scan_downsample", ikd->tree->h);
    resident[ikd->id] = scan.get();
    return true;
  }
  int sweep_points(const std::shared_ptr<LiDARData> &scan) const { return op.raw_total_points > 0 ? op.raw_total_points : scan->size(); }

  // LidarHelper::remove_motion_blur (LidarHelper.cpp:39-87) with the guard as intended (the reference's `if (get_interpolated_pose)
  // return false` at :54-58 is inverted and makes the function unreachable): point i was taken at
  // t_i = time + dt - (total - i) raw_point_dt and is moved into the frame of the scan time,
  //   T_LitoL = [R_GtoL R_GtoIi^T R_LtoI | R_GtoL (p_Ii + R_GtoIi^T p_LinI - p_LinG)]   (:76-81, fp32),
  // evaluated at `deskew_knots` times spread uniformly over the sweep; the device interpolates between the knots.
  bool deskew_knots(const std::shared_ptr<LiDARData> &scan, std::vector<float> &knots) {
    const int nk = std::max(2, op.deskew_knots), total = sweep_points(scan);
    const double dt = state->lidar_dt(scan->id), t1 = scan->time + dt, t0 = t1 - total * op.raw_point_dt;
    double RGtoL[9], pLinG[3];
    if (!LidarHelper::get_pose_LinG(*state, scan->id, t1, RGtoL, pLinG) || !LidarHelper::get_pose_LinG(*state, scan->id, t0, RGtoL, pLinG))
      return false;
    LidarHelper::get_pose_LinG(*state, scan->id, t1, RGtoL, pLinG);
    knots.resize((size_t)nk * 12);
    for (int j = 0; j < nk; j++) {
      double RGtoLi[9], pLiinG[3], RLitoG[9], R[9], d[3], p[3];
      if (!LidarHelper::get_pose_LinG(*state, scan->id, t0 + (t1 - t0) * j / (nk - 1), RGtoLi, pLiinG))
        return false;
      tr33(RGtoLi, RLitoG);
      mul33(RGtoL, RLitoG, R); // R_GtoL R_GtoLi^T = R_GtoL R_GtoIi^T R_LtoI
      for (int a = 0; a < 3; a++)
        d[a] = pLiinG[a] - pLinG[a];
      mv3(RGtoL, d, p);
      for (int a = 0; a < 3; a++) {
        for (int b = 0; b < 3; b++)
          knots[(size_t)j * 12 + 4 * a + b] = (float)R[3 * a + b];
        knots[(size_t)j * 12 + 4 * a + 3] = (float)p[a];
      }
    }
    return true;
  }

  // LidarHelper::transform_to_map (LidarHelper.cpp:115-194) from the CURRENT state: T_LtoM from the interpolated poses and,
  // with map_use_icp, the ICP block on the (uploaded, down-sampled) scan cloud. Leaves the scan resident, not yet transformed.
  bool transform_to_map(const std::shared_ptr<LiDARData> &scan, std::shared_ptr<iKDDATA> &ikd, float T[16], bool *icp_success) {
    const double dt = state->lidar_dt(scan->id);
    double RGtoL[9], pLinG[3], RGtoM[9], pMinG[3];
    if (!LidarHelper::get_pose_LinG(*state, scan->id, scan->time + dt, RGtoL, pLinG))
      return false; // "Cannot get pose for LiDAR" (:128-131)
    if (!LidarHelper::get_pose_LinG(*state, scan->id, ikd->time + dt, RGtoM, pMinG))
      return false;
    LidarHelper::T_LtoM(RGtoL, pLinG, RGtoM, pMinG, T);
    *icp_success = true;
    if (op.map_use_icp) {
      if (!upload(scan, ikd, op.raw_do_downsample))
        return false;
      *icp_success = run_icp(ikd, T);
    }
    return true;
  }
  // the PointMatcher<float>::ICP block (LidarHelper.cpp:156-187) on the resident scan cloud; T: initial guess in, lidar->T_LtoM out
  bool run_icp(std::shared_ptr<iKDDATA> &ikd, float T[16]) {
    ml_icp_options io;
    api.icp_default_options(&io);
    io.max_dist = op.map_icp_dist;
    io.seed = op.icp_seed++;
    ml_icp_result ir;
    const int irc = api.scan_icp(ikd->tree->h, T, &io, &ir);
    ck(irc, "ml_scan_icp", ikd->tree->h);
    if (irc == ML_OK)
      std::memcpy(T, ir.T_LtoM, sizeof(float) * 16);
    return irc == ML_OK && ir.success;
  }

  void erase_old(std::vector<std::shared_ptr<LiDARData>> &v) {
    for (auto it = v.begin(); it != v.end();) {
      if ((*it)->time + state->lidar_dt((*it)->id) < state->oldest_clone_time()) {
        pending.erase(it->get());
        it = v.erase(it);
      } else
        ++it;
    }
  }
  // UpdaterLidar::erase_invalid_dataset (UpdaterLidar.cpp:127-145)
  void erase_invalid_dataset() {
    erase_old(stack_lidar_raw);
    erase_old(stack_lidar_new);
    erase_old(stack_lidar_used);
  }

  // LidarHelper::register_scan (LidarHelper.cpp:196-207): "If ICP failed, retry it here" with the state the update has just
  // improved; the scan is added to the map - and the 1 s staleness clock (UpdaterLidar.cpp:91) reset - only if the alignment
  // succeeded. A scan whose ICP fails twice is dropped, so a bad T_LtoM never reaches the map.
  void register_scan(const std::shared_ptr<LiDARData> &scan, std::shared_ptr<iKDDATA> &ikd) {
    auto p = pending.find(scan.get());
    if (p == pending.end())
      return;
    Pending &pd = p->second;
    if (!pd.icp_success) {
      float T[16];
      bool ok = false;
      if (transform_to_map(scan, ikd, T, &ok)) {
        std::memcpy(pd.T_LtoM, T, sizeof(T));
        pd.icp_success = ok;
        if (resident[ikd->id] == scan.get()) // uploaded for the ICP: pointcloud_in_map with the new transform
          ck(api.scan_transform(ikd->tree->h, pd.T_LtoM), "ml_scan_transform", ikd->tree->h);
      }
    }
    if (pd.icp_success) {
      if (resident[ikd->id] != scan.get()) { // another scan of this LiDAR was processed in between: bring this one back
        if (upload(scan, ikd, false))
          ck(api.scan_transform(ikd->tree->h, pd.T_LtoM), "ml_scan_transform", ikd->tree->h);
      }
      if (resident[ikd->id] == scan.get()) {
        ck(api.map_add_scan(ikd->tree->h, op.map_do_downsample ? 1 : 0, (float)op.map_downsample_size, nullptr), "ml_map_add_scan", ikd->tree->h);
        ikd->last_up_time = scan->time;
      }
    }
    pending.erase(p);
  }

  // LidarHelper::propagate_map_to_newest_clone (LidarHelper.cpp:209-266): pose look-ups here, point work on the device
  void propagate_map_to_newest_clone(std::shared_ptr<iKDDATA> &ikd) {
    const double dt = state->lidar_dt(ikd->id);
    double RGtoLold[9], pLoldinG[3], RGtoLnew[9], pLnewinG[3];
    if (!LidarHelper::get_pose_LinG(*state, ikd->id, ikd->time + dt, RGtoLold, pLoldinG)) {
      std::fprintf(stderr, "\033[31m[LiDAR-GPU]propagate_map_to_newest_clone::Cannot get the bounding poses for Map %d (t=%.4f)\033[0m\n",
                   ikd->id, ikd->time + dt);
      std::exit(EXIT_FAILURE); // LidarHelper.cpp:244
    }
    LidarHelper::get_pose_LinG(*state, ikd->id, state->newest_clone_time(), RGtoLnew, pLnewinG);
    double RLoldtoG[9], R[9], d[3], p[3];
    tr33(RGtoLold, RLoldtoG);
    mul33(RGtoLnew, RLoldtoG, R);
    for (int i = 0; i < 3; i++)
      d[i] = pLoldinG[i] - pLnewinG[i];
    mv3(RGtoLnew, d, p);
    double T[16] = {R[0], R[1], R[2], p[0], R[3], R[4], R[5], p[1], R[6], R[7], R[8], p[2], 0, 0, 0, 1};
    ck(api.map_reanchor(ikd->tree->h, T, state->time() - FT, op.map_decay_time, op.map_decay_dist, op.map_do_downsample ? 1 : 0,
                        (float)op.map_downsample_size, nullptr),
       "ml_map_reanchor", ikd->tree->h);
    ikd->time = state->newest_clone_time() - dt; // init_map_local(prop = true) keeps last_up_time
  }

  // UpdaterLidar::update (UpdaterLidar.cpp:163-408)
  bool update(const std::shared_ptr<LiDARData> &lidar, std::shared_ptr<iKDDATA> &ikd, UpdaterStatistics &chi) {
    // ---- states involved: H_order / map_hx (UpdaterLidar.cpp:171-214, StateHelper::insert_map) ----
    std::vector<const void *> H_order;
    std::vector<int> widths;
    std::map<const void *, int> map_hx;
    int total_hx = 0;
    auto insert_map = [&](const void *var, int width) {
      if (map_hx.find(var) == map_hx.end()) {
        map_hx[var] = total_hx;
        H_order.push_back(var);
        widths.push_back(width);
        total_hx += width;
      }
    };
    const double dt = state->lidar_dt(lidar->id);
    if (op.do_calib_ext)
      insert_map(state->calib_var(lidar->id), 6);
    if (op.do_calib_dt)
      insert_map(state->dt_var(lidar->id), 1);
    double RGtoI_est[9], RGtoA_est[9], RGtoI_fej[9], RGtoA_fej[9], pIinG_est[3], pAinG_est[3], pIinG_fej[3], pAinG_fej[3];
    std::vector<JacBlock> dTdxI, dTdxA;
    if (!state->get_interpolated_jacobian(lidar->time + dt, lidar->id, RGtoI_fej, pIinG_fej, dTdxI))
      return false;
    for (auto &b : dTdxI)
      insert_map(b.var, b.width);
    if (!state->get_interpolated_jacobian(ikd->time + dt, lidar->id, RGtoA_fej, pAinG_fej, dTdxA)) {
      if (ikd->time + dt > state->oldest_clone_time())
        return false;
      std::fprintf(stderr, "\033[31m[LiDAR-GPU]get_linsys::Cannot get the bounding poses for Map %d\033[0m\n", lidar->id);
      std::exit(EXIT_FAILURE); // UpdaterLidar.cpp:210
    }
    for (auto &b : dTdxA)
      insert_map(b.var, b.width);
    if (total_hx > ML_MAX_HX || dTdxI.size() > ML_MAX_BLOCKS || dTdxA.size() > ML_MAX_BLOCKS)
      die("update: state blocks exceed ML_MAX_HX / ML_MAX_BLOCKS", nullptr);
    state->get_interpolated_pose(lidar->time + dt, RGtoI_est, pIinG_est);
    state->get_interpolated_pose(ikd->time + dt, RGtoA_est, pAinG_est);

    // ---- precomputed transformations (UpdaterLidar.cpp:223-254) -> ml_scan_consts ----
    ml_scan_consts c;
    std::memset(&c, 0, sizeof(c));
    double RItoL[9], RLtoI[9], pIinL[3], pLinI[3], t3[3], d3[3];
    state->calibration(lidar->id, RItoL, pIinL);
    tr33(RItoL, RLtoI);
    mv3(RLtoI, pIinL, pLinI);
    for (int i = 0; i < 3; i++)
      pLinI[i] = -pLinI[i];
    auto lidar_frame = [&](const double *RGtoX, const double *pXinG, double *RGtoLx, double *pLxinG) {
      double RXtoG[9];
      mul33(RItoL, RGtoX, RGtoLx);
      tr33(RGtoX, RXtoG);
      mv3(RXtoG, pLinI, t3);
      for (int i = 0; i < 3; i++)
        pLxinG[i] = pXinG[i] + t3[i];
    };
    double RGtoL_est[9], pLinG_est[3], RGtoM_est[9], pMinG_est[3], RGtoL_fej[9], pLinG_fej[3], pMinG_fej[3], RLtoG[9];
    lidar_frame(RGtoI_est, pIinG_est, RGtoL_est, pLinG_est);
    lidar_frame(RGtoA_est, pAinG_est, RGtoM_est, pMinG_est);
    lidar_frame(RGtoI_fej, pIinG_fej, RGtoL_fej, pLinG_fej);
    lidar_frame(RGtoA_fej, pAinG_fej, c.RGtoM_fej, pMinG_fej);
    tr33(RGtoL_est, RLtoG);
    mul33(RGtoM_est, RLtoG, c.RLtoM_est);
    for (int i = 0; i < 3; i++)
      d3[i] = pLinG_est[i] - pMinG_est[i];
    mv3(RGtoM_est, d3, c.pLinM_est);
    tr33(RGtoL_fej, c.RLtoG_fej);
    mul33(c.RGtoM_fej, c.RLtoG_fej, c.RLtoM_fej);
    for (int i = 0; i < 3; i++)
      d3[i] = pLinG_fej[i] - pMinG_fej[i];
    mv3(c.RGtoM_fej, d3, c.pLinM_fej);
    tr33(RGtoI_fej, c.RItoG_fej);
    std::memcpy(c.RLtoI_fej, RLtoI, sizeof(RLtoI));
    std::memcpy(c.RItoL_fej, RItoL, sizeof(RItoL));
    std::memcpy(c.RGtoA_fej, RGtoA_fej, sizeof(RGtoA_fej));
    std::memcpy(c.pIinL_fej, pIinL, sizeof(pIinL));
    std::memcpy(c.pLinI_fej, pLinI, sizeof(pLinI));
    std::memcpy(c.pIinG_fej, pIinG_fej, sizeof(pIinG_fej));
    std::memcpy(c.pAinG_fej, pAinG_fej, sizeof(pAinG_fej));
    c.nI = (int)dTdxI.size(), c.nA = (int)dTdxA.size();
    for (int j = 0; j < c.nI; j++) {
      c.dTdxI[j].col = map_hx.at(dTdxI[j].var), c.dTdxI[j].width = dTdxI[j].width;
      std::memcpy(c.dTdxI[j].D, dTdxI[j].D, sizeof(c.dTdxI[j].D));
    }
    for (int j = 0; j < c.nA; j++) {
      c.dTdxA[j].col = map_hx.at(dTdxA[j].var), c.dTdxA[j].width = dTdxA[j].width;
      std::memcpy(c.dTdxA[j].D, dTdxA[j].D, sizeof(c.dTdxA[j].D));
    }
    c.calib_col = op.do_calib_ext ? map_hx.at(state->calib_var(lidar->id)) : -1;
    c.at_clone_m = state->have_clone(lidar->time + dt) ? 1 : 0;
    c.at_clone_a = state->have_clone(ikd->time + dt) ? 1 : 0;
    c.ie_o = state->intr_ori_cov(), c.ie_p = state->intr_pos_cov();
    c.total_hx = total_hx;
    c.raw_noise = op.raw_noise, c.map_noise = op.map_noise;
    c.map_ngbr_num = op.map_ngbr_num, c.map_ngbr_max_d = op.map_ngbr_max_d;
    c.plane_max_p2pd = op.plane_max_p2pd, c.plane_max_condi = op.plane_max_condi;
    c.chi2_mult = op.chi2_mult; // the library applies quantile(chi_squared(k - 2), 0.95), UpdaterStatistics.cpp:31-33,110-111

    // ---- marginal covariance for the chi2 test (UpdaterLidar.cpp:257) ----
    std::vector<double> P((size_t)total_hx * total_hx);
    state->get_marginal_covariance(H_order, widths, P.data());

    // ---- transform_to_map (LidarHelper.cpp:141-151) + the device pipeline: one call, one synchronisation ----
    Pending pend;
    LidarHelper::T_LtoM(RGtoL_est, pLinG_est, RGtoM_est, pMinG_est, pend.T_LtoM);
    std::vector<double> HtH((size_t)total_hx * total_hx), Htr(total_hx);
    ml_counters cnt;
    std::memset(&cnt, 0, sizeof(cnt));
    int rc;
    if (!op.map_use_icp && !op.raw_remove_motion_blur) {
      rc = api.scan_update(ikd->tree->h, lidar->xyz, 4, lidar->size(), (float)(lidar->time - FT), (float)op.min_range,
                           (float)op.max_range, op.raw_do_downsample ? (float)op.raw_downsample_size : 0.f, pend.T_LtoM, &c,
                           P.data(), HtH.data(), Htr.data(), &cnt);
    } else {
      // the separate calls: LiDARData ctor, [remove_motion_blur] (UpdaterLidar.cpp:82), [downsample] (:88), then
      // transform_to_map with the ICP block (LidarHelper.cpp:156-187): the scan cloud is matched against the map starting
      // from the state's T_LtoM; on success the ICP transform replaces it for the data association AND for register_scan
      // (lidar->T_LtoM), while the residual/Jacobian keep using the state estimate (UpdaterLidar.cpp:281-283).
      if (!upload(lidar, ikd, op.raw_do_downsample))
        return false; // de-skew poses not available yet: keep the scan (UpdaterLidar.cpp:82-85)
      if (op.map_use_icp)
        pend.icp_success = run_icp(ikd, pend.T_LtoM); // lidar->icp_success
      last_icp_success = pend.icp_success;
      ck(api.scan_transform(ikd->tree->h, pend.T_LtoM), "ml_scan_transform", ikd->tree->h);
      rc = api.linsys(ikd->tree->h, &c, P.data(), HtH.data(), Htr.data(), &cnt);
    }
    ck(rc, "ml_scan_update", ikd->tree->h);
    resident[ikd->id] = lidar.get();
    last_counters = cnt;
    chi.add(cnt);
    if (rc == ML_SOFT_NO_ROWS || rc == ML_SOFT_NO_MAP)
      return false; // H_big.rows() < 1 (UpdaterLidar.cpp:398-399)

    // ---- measurement compression stand-in + the untouched EKF update (UpdaterLidar.cpp:397-406) ----
    std::vector<double> Rc((size_t)total_hx * total_hx), rc_((size_t)total_hx);
    const int rank = api.info_to_sqrt(HtH.data(), Htr.data(), total_hx, Rc.data(), rc_.data());
    if (rank < 1)
      return false;
    state->EKFUpdate(H_order, widths, Rc.data(), rc_.data(), rank, total_hx);
    pending[lidar.get()] = pend;
    n_updates++;
    return true;
  }

  StateView *state;
  ml_api api;
  int cap_scan;
  std::vector<float *> pool; // free pinned scan buffers (ml_host_alloc), each max_scan_points PointXYZ
};

} // namespace mins_gpu
