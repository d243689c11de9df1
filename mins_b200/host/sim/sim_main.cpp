// mins_sim_run: one closed-loop IMU + LiDAR simulation run in host C++ (see harness.hpp). Prints one JSON line.
//   mins_sim_run --bspline F --floorplan F [--seed S | --seeds S0,S1,...] [--duration T] [--icp 0|1] [--no-lidar] [--no-noise] [--no-gt-map]
//                [--device D] [--traj out.txt] [--state-log out.txt] [--deskew 0|1]
// Without --no-lidar a CUDA device is required (the LiDAR updater has no CPU path); --no-lidar runs the IMU + clone-window core
// alone (the CPU tests use it to pin this file against the Python harness it replaces).
#include "harness.hpp"

#include <cinttypes>

using namespace mins_sim;

static SimResult run(const SimConfig &cfg) {
  using clk = std::chrono::steady_clock;
  const auto t_begin = clk::now();
  SimResult res;
  BsplineSE3 spline(load_bspline_dataset(cfg.bspline_path));
  std::mt19937_64 rng_imu(cfg.seed * 7919ull + 1), rng_init(cfg.seed * 7919ull + 3);
  std::normal_distribution<double> nd(0.0, 1.0);
  const double nz = cfg.no_noise ? 0.0 : 1.0;

  // lidar0 extrinsics of config/simulation/config_lidar.yaml:33-37: T_imu_lidar = [R_LtoI | p_LinI], a 15 deg roll
  const double c15 = std::cos(M_PI / 12.0), s15 = std::sin(M_PI / 12.0);
  const double R_LtoI[9] = {1, 0, 0, 0, c15, -s15, 0, s15, c15}, p_LinI[3] = {0.1, 0.1, 0.1};
  double R_ItoL[9], p_IinL[3];
  m3_T(R_LtoI, R_ItoL);
  m3_vec(R_ItoL, p_LinI, p_IinL);
  for (int i = 0; i < 3; i++)
    p_IinL[i] = -p_IinL[i];

  const double dt_imu = 1.0 / cfg.freq_imu;
  const double t0 = std::ceil((spline.t_start + 0.1) * cfg.clone_freq) / cfg.clone_freq;
  double R0[9], p0[3], w0[3], v0[3], a0[3];
  if (!spline.get_kinematics(t0, R0, p0, w0, v0, a0))
    throw std::runtime_error("trajectory too short");
  State st;
  {
    double e[3], E[9];
    for (int i = 0; i < 3; i++)
      e[i] = nz * cfg.init_ori_err * nd(rng_init);
    exp_so3(e, E);
    m3_mul(E, R0, st.R);
    for (int i = 0; i < 3; i++) {
      st.p[i] = p0[i] + nz * cfg.init_pos_err * nd(rng_init);
      st.v[i] = v0[i], st.bg[i] = 0, st.ba[i] = 0;
    }
    st.P = Mat(15, 15);
    for (int i = 0; i < 3; i++) {
      st.P(i, i) = cfg.init_ori_err * cfg.init_ori_err, st.P(3 + i, 3 + i) = cfg.init_pos_err * cfg.init_pos_err;
      st.P(6 + i, 6 + i) = 0.01 * 0.01, st.P(9 + i, 9 + i) = 1e-4 * 1e-4 + 1e-8, st.P(12 + i, 12 + i) = 0.02 * 0.02;
    }
    st.intr_order = cfg.intr_order;
    std::memcpy(st.R_ItoL, R_ItoL, sizeof(R_ItoL));
    std::memcpy(st.p_IinL, p_IinL, sizeof(p_IinL));
    st.time = t0;
  }
  SimStateView view(&st);
  const double noises[4] = {cfg.gyro_noise, cfg.accel_noise, cfg.gyro_walk, cfg.accel_walk};
  double bg_true[3] = {0, 0, 0}, ba_true[3] = {0, 0, 0};
  auto imu_meas = [&](double t, double *wm, double *am) {
    double R[9], p[3], w[3], v[3], a[3], ag[3], Ra[3];
    if (!spline.get_kinematics(t, R, p, w, v, a))
      throw std::runtime_error("IMU time outside the spline");
    const double sd = 1.0 / std::sqrt(dt_imu);
    for (int i = 0; i < 3; i++)
      ag[i] = a[i] + st.g[i];
    m3_vec(R, ag, Ra);
    for (int i = 0; i < 3; i++) {
      wm[i] = w[i] + bg_true[i] + nz * cfg.gyro_noise * sd * nd(rng_imu);
      am[i] = Ra[i] + ba_true[i] + nz * cfg.accel_noise * sd * nd(rng_imu);
    }
    for (int i = 0; i < 3; i++) {
      bg_true[i] += nz * cfg.gyro_walk * std::sqrt(dt_imu) * nd(rng_imu);
      ba_true[i] += nz * cfg.accel_walk * std::sqrt(dt_imu) * nd(rng_imu);
    }
  };

  // ---- LiDAR: the updater (C++ shim over the C ABI) and the device-side ray-caster -------------------------------------
  std::unique_ptr<mins_gpu::UpdaterLidar> upd;
  mins_gpu::ml_api api = mins_gpu::ml_api::real();
  ml_handle *caster = nullptr;
  std::vector<double> dirs_L;
  std::vector<float> scan4;
  const double v_angles_arr[17] = {-80, -70, -60, -50, -40, -30, -20, -10, 0, 10, 20, 30, 40, 50, 60, 70, 80}; // config_lidar.yaml:38
  std::vector<double> h_angles;
  for (double h = -180.0; h <= 180.0 + 1e-9; h += cfg.h_resolution)
    h_angles.push_back(h);
  auto raycast = [&](double t, const std::vector<double> &dirs, unsigned long long seed, int *n_out) {
    double R_gt[9], p_gt[3], RIG[9], pL[3], R_LtoG[9], tmp[3];
    spline.get_pose(t, R_gt, p_gt);
    m3_T(R_gt, RIG);
    m3_vec(RIG, p_LinI, tmp);
    for (int i = 0; i < 3; i++)
      pL[i] = p_gt[i] + tmp[i];
    m3_mul(RIG, R_LtoI, R_LtoG);
    const int n_rays = (int)(dirs.size() / 3);
    if (ml_sim_set_rays(caster, dirs.data(), n_rays) != ML_OK)
      throw std::runtime_error(std::string("ml_sim_set_rays: ") + ml_last_error(caster));
    scan4.resize((size_t)4 * n_rays);
    if (ml_sim_raycast(caster, pL, R_LtoG, cfg.op.min_range, cfg.op.max_range, cfg.no_noise ? 0.f : (float)cfg.op.raw_noise, seed,
                       scan4.data(), n_rays, n_out) != ML_OK)
      throw std::runtime_error(std::string("ml_sim_raycast: ") + ml_last_error(caster));
  };
  if (cfg.use_lidar) {
    upd.reset(new mins_gpu::UpdaterLidar(&view, cfg.op, api, 1 << 17, 1 << 19, cfg.device));
    ml_config cc{cfg.device, 1 << 17, 1024, 0, 0.f};
    if (ml_create(&cc, &caster) != ML_OK)
      throw std::runtime_error(std::string("ml_create (ray-caster): ") + ml_global_error());
    const std::vector<double> planes = load_floorplan_planes(cfg.floorplan_path);
    if (ml_sim_set_world(caster, planes.data(), (int)(planes.size() / 22)) != ML_OK)
      throw std::runtime_error("ml_sim_set_world failed");
    dirs_L = ray_dirs(std::vector<double>(v_angles_arr, v_angles_arr + 17), h_angles);
    if (cfg.use_gt_lidar) {
      // Initializer::init_lidar_sim (mins/src/init/Initializer.cpp:308-340): v_angles = -90..89; the 180 values -90..89 are
      // APPENDED to the h_angles the options already hold: 180 x 361 rays; the scan at the true pose becomes the map as it is
      std::vector<double> vd, hd = h_angles;
      for (int j = -90; j < 90; j++) {
        vd.push_back(j);
        hd.push_back(j);
      }
      int n = 0;
      raycast(t0, ray_dirs(vd, hd), (cfg.seed << 32) | 0xffffffffull, &n);
      upd->init_map_gt(0, t0, scan4.data(), n, 4);
    }
  }

  const int n_steps = (int)std::llround(std::min(cfg.duration, spline.t_end - 0.05 - t0) * cfg.freq_imu);
  const int imu_per_clone = (int)std::llround(cfg.freq_imu / cfg.clone_freq), imu_per_scan = (int)std::llround(cfg.freq_imu / cfg.freq_lidar);
  std::vector<PoseSample> est, gt;
  double nees_sum = 0;
  double prev_w[3], prev_a[3], cur_w[3], cur_a[3];
  imu_meas(t0, prev_w, prev_a);
  st.augment_clone();
  for (int k = 1; k <= n_steps; k++) {
    const double t = t0 + k * dt_imu;
    imu_meas(t, cur_w, cur_a);
    double wm[3], am[3];
    for (int i = 0; i < 3; i++)
      wm[i] = 0.5 * (prev_w[i] + cur_w[i]), am[i] = 0.5 * (prev_a[i] + cur_a[i]);
    st.propagate(wm, am, dt_imu, noises);
    st.time = t; // keep clone keys on the exact grid
    std::memcpy(prev_w, cur_w, sizeof(cur_w)), std::memcpy(prev_a, cur_a, sizeof(cur_a));
    if (k % imu_per_clone == 0) { // SystemManager::feed_measurement_imu: clone, marginalise, re-anchor the map (:99-105)
      st.augment_clone();
      while (st.newest_clone_time() - st.oldest_clone_time() > cfg.window_size + 1e-9)
        st.marginalize_oldest();
      if (upd) {
        const auto c0 = clk::now();
        upd->propagate_map_frame();
        res.ldr_wall_s += std::chrono::duration<double>(clk::now() - c0).count();
      }
      PoseSample e, g;
      e.t = g.t = t;
      std::memcpy(e.R, st.R, sizeof(e.R)), std::memcpy(e.p, st.p, sizeof(e.p));
      spline.get_pose(t, g.R, g.p);
      {
        int q = 0;
        for (int blk = 0; blk < 2; blk++) // orientation block, then position block: entries 00 01 02 11 12 22
          for (int a = 0; a < 3; a++)
            for (int b = a; b < 3; b++)
              e.cov[q++] = st.P(3 * blk + a, 3 * blk + b);
      }
      est.push_back(e), gt.push_back(g);
      double Pp[9], Pi[9], d[3], Pd[3];
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++)
          Pp[3 * a + b] = st.P(3 + a, 3 + b);
      for (int a = 0; a < 3; a++)
        d[a] = st.p[a] - g.p[a];
      if (m3_inv(Pp, Pi)) {
        m3_vec(Pi, d, Pd);
        nees_sum += d[0] * Pd[0] + d[1] * Pd[1] + d[2] * Pd[2];
      }
    }
    if (upd && k % imu_per_scan == 0) { // SystemManager::feed_measurement_lidar (:177-185)
      int n = 0;
      const auto r0 = clk::now();
      raycast(t, dirs_L, (cfg.seed << 32) | (unsigned long long)(k / imu_per_scan), &n);
      const auto c0 = clk::now();
      res.raycast_wall_s += std::chrono::duration<double>(c0 - r0).count();
      upd->feed_measurement(0, t, scan4.data(), n, 4);
      upd->try_update();
      res.ldr_wall_s += std::chrono::duration<double>(clk::now() - c0).count();
      res.n_scans++;
    }
  }
  ate(est, gt, &res.rmse_ori_deg, &res.rmse_pos_m);
  res.nees_pos = nees_sum / (double)est.size() / 3.0;
  res.n_clones = (int)est.size();
  {
    double d[3];
    for (int a = 0; a < 3; a++)
      d[a] = est.back().p[a] - gt.back().p[a];
    res.final_pos_err = norm3(d);
  }
  for (auto &e : est) {
    double q[4];
    rot_to_quat_jpl(e.R, q);
    std::vector<double> row = {e.t, e.p[0], e.p[1], e.p[2], q[0], q[1], q[2], q[3]};
    row.insert(row.end(), e.cov, e.cov + 12);
    res.traj.push_back(row);
  }
  if (upd) {
    res.n_updates = upd->n_updates;
    res.n_acc = upd->Chi[0].n_acp, res.n_meas = upd->Chi[0].n_meas;
    res.map_points = upd->ikd_data[0]->tree->size();
    upd.reset();
    ml_destroy(caster);
  }
  res.wall_s = std::chrono::duration<double>(clk::now() - t_begin).count();
  return res;
}

int main(int argc, char **argv) {
  SimConfig cfg;
  std::vector<unsigned long long> seeds;
  // simulation option set (mins/config/simulation/config_lidar.yaml:11-29)
  cfg.op.max_n = 1, cfg.op.max_range = 100.0, cfg.op.min_range = 1.0, cfg.op.chi2_mult = 1.0;
  cfg.op.raw_do_downsample = true, cfg.op.raw_downsample_size = 0.3, cfg.op.raw_noise = 0.01;
  cfg.op.map_do_downsample = true, cfg.op.map_downsample_size = 0.3, cfg.op.map_noise = 0.5, cfg.op.map_ngbr_num = 5;
  cfg.op.map_ngbr_max_d = 1.0, cfg.op.map_decay_time = 120, cfg.op.map_decay_dist = 100, cfg.op.map_icp_dist = 50, cfg.op.map_use_icp = false;
  cfg.op.plane_max_p2pd = 0.1, cfg.op.plane_max_condi = 200.0;
  for (int i = 1; i < argc; i++) {
    const std::string a = argv[i];
    auto next = [&]() -> const char * {
      if (i + 1 >= argc) {
        std::fprintf(stderr, "missing value for %s\n", a.c_str());
        std::exit(2);
      }
      return argv[++i];
    };
    if (a == "--bspline")
      cfg.bspline_path = next();
    else if (a == "--floorplan")
      cfg.floorplan_path = next();
    else if (a == "--seed")
      cfg.seed = strtoull(next(), nullptr, 10);
    else if (a == "--seeds") { // several runs in ONE process (one CUDA context): a Monte-Carlo worker
      std::string list = next();
      for (size_t pos = 0; pos < list.size();) {
        const size_t comma = list.find(',', pos);
        seeds.push_back(strtoull(list.substr(pos, comma - pos).c_str(), nullptr, 10));
        pos = comma == std::string::npos ? list.size() : comma + 1;
      }
    }
    else if (a == "--duration")
      cfg.duration = atof(next());
    else if (a == "--icp")
      cfg.op.map_use_icp = atoi(next()) != 0;
    else if (a == "--deskew")
      cfg.op.raw_remove_motion_blur = atoi(next()) != 0;
    else if (a == "--device")
      cfg.device = atoi(next());
    else if (a == "--traj")
      cfg.traj_out = next();
    else if (a == "--state-log") // the estimate in State_Logger::save_trajectory_to_file's format (what mins_eval reads)
      cfg.state_log_out = next();
    else if (a == "--no-lidar")
      cfg.use_lidar = false;
    else if (a == "--no-noise")
      cfg.no_noise = true;
    else if (a == "--no-gt-map")
      cfg.use_gt_lidar = false;
    else {
      std::fprintf(stderr, "unknown argument %s\n", a.c_str());
      return 2;
    }
  }
  if (cfg.bspline_path.empty() || (cfg.use_lidar && cfg.floorplan_path.empty())) {
    std::fprintf(stderr, "usage: mins_sim_run --bspline F --floorplan F [--seed S] [--duration T] [--icp 0|1] [--no-lidar] [--no-noise]\n");
    return 2;
  }
  if (seeds.empty())
    seeds.push_back(cfg.seed);
  for (unsigned long long seed : seeds)
  try {
    cfg.seed = seed;
    cfg.op.icp_seed = cfg.seed << 20;
    const SimResult r = run(cfg);
    if (!cfg.traj_out.empty()) {
      FILE *f = std::fopen(cfg.traj_out.c_str(), "w");
      for (auto &row : r.traj)
        std::fprintf(f, "%.9f %.12g %.12g %.12g %.12g %.12g %.12g %.12g\n", row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]);
      std::fclose(f);
    }
    if (!cfg.state_log_out.empty()) {
      // mins/src/utils/State_Logger.h:323-340: time px py pz qx qy qz qw (6 significant digits), then the upper triangles of the
      // orientation and the position covariance blocks (10 significant digits): the file mins_eval's run_comparison loads
      FILE *f = std::fopen(cfg.state_log_out.c_str(), "w");
      for (auto &row : r.traj) {
        for (int c = 0; c < 8; c++)
          std::fprintf(f, "%.6g ", row[c]);
        for (int c = 8; c < 20; c++)
          std::fprintf(f, c < 19 ? "%.10g " : "%.10g\n", row[c]);
      }
      std::fclose(f);
    }
    std::printf("{\"seed\": %llu, \"icp\": %d, \"rmse_ori_deg\": %.9g, \"rmse_pos_m\": %.9g, \"nees_pos\": %.6g, \"final_pos_err\": %.6g, "
                "\"n_updates\": %d, \"n_scans\": %d, \"n_clones\": %d, \"n_acc\": %ld, \"n_meas\": %ld, \"map_points\": %d, "
                "\"wall_s\": %.4f, \"ldr_wall_s\": %.4f, \"raycast_wall_s\": %.4f}\n",
                cfg.seed, cfg.op.map_use_icp ? 1 : 0, r.rmse_ori_deg, r.rmse_pos_m, r.nees_pos, r.final_pos_err, r.n_updates, r.n_scans,
                r.n_clones, r.n_acc, r.n_meas, r.map_points, r.wall_s, r.ldr_wall_s, r.raycast_wall_s);
    std::fflush(stdout);
  } catch (const std::exception &e) {
    std::fprintf(stderr, "mins_sim_run: %s\n", e.what());
    return 1;
  }
  return 0;
}
