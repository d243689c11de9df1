/*
 * mins_lidar.h — C ABI of the B200-native LiDAR measurement-update path for rpng/MINS.
 *
 * One opaque handle = one LiDAR id = one local map (the reference keeps one `iKDDATA` per id,
 * mins/src/update/lidar/UpdaterLidar.cpp:43-47). Every entry point names the reference interface it replaces.
 * Plain pointers and sizes only; the caller owns all host buffers, the library owns all device memory
 * (pre-sized at ml_create, nothing is allocated on the per-scan path).
 *
 * Status convention (reference error behaviour):
 *    0  ok
 *   >0  soft: "cannot use this scan" (the caller returns false and retries/drops it, as UpdaterLidar.cpp:398-399)
 *   <0  fatal: CUDA error / capacity exceeded / bad argument (the C++ shim prints and exits like UpdaterLidar.cpp:210)
 * No C++ exception crosses this boundary. ml_last_error() gives the text of the last non-zero status.
 *
 * Threading: one host thread per handle, one CUDA stream per handle; different handles are independent.
 * There is NO CPU fallback: if no CUDA device is usable ml_create fails with ML_ERR_CUDA.
 *
 * Point layout: float[4] = {x, y, z, intensity}; intensity carries "seconds since the first scan (FT)" exactly as
 * LiDARData's constructor stores it (mins/src/update/lidar/LidarTypes.cpp:47).
 */
#ifndef MINS_LIDAR_H
#define MINS_LIDAR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ML_MAX_HX 64     /* max columns of the measurement Jacobian (reference: 48, 54 or 55) */
#define ML_MAX_BLOCKS 8  /* max Jacobian blocks per interpolated pose (intr_order + 1 poses + time offset) */
#define ML_MAX_K 10      /* max map_ngbr_num */

#define ML_OK 0
#define ML_SOFT_NO_ROWS 1   /* no point passed the gates: UpdaterLidar::update returns false (UpdaterLidar.cpp:398) */
#define ML_SOFT_NO_MAP 2    /* map not initialised (tree->initialized() false, UpdaterLidar.cpp:91) */
#define ML_ERR_CUDA (-1)
#define ML_ERR_CAPACITY (-2)
#define ML_ERR_ARG (-3)

typedef struct ml_handle ml_handle;

typedef struct ml_config {
  int device;          /* CUDA device ordinal */
  int max_scan_points; /* capacity for one raw scan */
  int max_map_points;  /* capacity of the local map */
  int max_grid_cells;  /* cap on cells of the dense kNN grid; 0 = default (1<<24) */
  float grid_cell;     /* kNN grid cell edge [m]; 0 = default 0.5 */
} ml_config;

/* One block of d(interpolated pose)/d(state variable): the reference's `dTdx[j]` with `map_hx.at(order[j])`
 * (mins/src/state/State.cpp:745-816, UpdaterLidar.cpp:343-347). D is 6 x width, row-major with row stride 6. */
typedef struct ml_jac_block {
  int col;   /* first column in H (map_hx) */
  int width; /* 6 for a pose, 1 for the time offset */
  double D[36];
} ml_jac_block;

/* Every per-scan constant UpdaterLidar::update computes before its point loop (UpdaterLidar.cpp:171-257).
 * 3x3 matrices are row-major. Names follow the reference's local variables. */
typedef struct ml_scan_consts {
  double RLtoM_est[9], pLinM_est[3]; /* :246-247 */
  double RLtoM_fej[9], pLinM_fej[3]; /* :248-249 */
  double RGtoM_fej[9];               /* :243 */
  double RItoG_fej[9];               /* :236 */
  double RLtoI_fej[9], RItoL_fej[9]; /* :227-228 */
  double RGtoA_fej[9];               /* :197 */
  double RLtoG_fej[9];               /* :238 */
  double pIinL_fej[3];               /* :229 */
  double pLinI_fej[3];               /* :230 */
  double pIinG_fej[3], pAinG_fej[3]; /* :190,:197 */
  int nI, nA;                        /* dTdxI.size(), dTdxA.size() */
  ml_jac_block dTdxI[ML_MAX_BLOCKS]; /* :190 */
  ml_jac_block dTdxA[ML_MAX_BLOCKS]; /* :197 */
  int calib_col;                     /* map_hx.at(calibration) if do_calib_ext else -1 (:181,:349) */
  int at_clone_m, at_clone_a;        /* :251-252 */
  double ie_o, ie_p;                 /* :253-254 */
  int total_hx;                      /* :172 */
  /* OptionsLidar (mins/src/options/OptionsLidar.h:38-99) */
  double raw_noise, map_noise;
  int map_ngbr_num;
  double map_ngbr_max_d; /* compared against the SQUARED distance of the k-th neighbour (LidarHelper.cpp:284) */
  double plane_max_p2pd;
  double plane_max_condi;
  double chi2_mult; /* OptionsLidar::chi2_mult. The gate is chi2 < chi2_mult * quantile(chi_squared(res.rows()), 0.95) with
                       res.rows() = map_ngbr_num + 1 - 3 after the nullspace projection (UpdaterLidar.cpp:301,362,
                       UpdaterStatistics.cpp:31-33,110-111): 3 DoF (7.814727903251179) at the default k = 5. The library holds
                       the quantile table for k = 3..ML_MAX_K itself. */
} ml_scan_consts;

/* What UpdaterStatistics would have accumulated (mins/src/update/UpdaterStatistics.cpp:47-83), as order-independent
 * moments, plus why points were dropped. */
typedef struct ml_counters {
  int n_points;     /* points in the (down-sampled) scan */
  int n_knn_fail;   /* LidarHelper::get_neighbors false (LidarHelper.cpp:280-285) */
  int n_plane_fail; /* LidarHelper::compute_plane false (LidarHelper.cpp:310,322) */
  int n_badmatch;   /* UpdaterLidar.cpp:296 */
  int n_chi_rej;    /* Chi2Check false (incl. NaN) */
  int n_acc;        /* accepted points (3 rows each in the reference's H_big) */
  double sum_chi, sum_chi2; /* over all points that reached the chi2 test */
  double sum_res, sum_res2; /* over the k - 2 projected residuals of those points: sum of r' (filled only with residual statistics on,
                               ml_set_stats: it depends on the reference's Givens basis, which costs a third of the rows kernel on
                               top; 0 otherwise) and sum of r'^2 (always) */
} ml_counters;

/* per-point debug/parity export (ml_point_rows) */
typedef struct ml_point_row {
  int status;     /* 0 accepted, 1 knn fail, 2 plane fail, 3 bad match, 4 chi2 reject */
  int pad;
  double plane[4]; /* [abc, d] un-normalised (LidarHelper.cpp:315-316) */
  double res;      /* whitened scan-point residual, row 5 (UpdaterLidar.cpp:313,358) */
  double chi2;     /* UpdaterStatistics.cpp:110 */
  double w;        /* ||v||^2 : weight of h h^T in H'^T H' */
  double vr;       /* v^T r' : H'^T r' = h * vr */
  double raw_n;    /* UpdaterLidar.cpp:353 */
} ml_point_row;

/* ---- lifetime -------------------------------------------------------------------------------------------- */
/* replaces: iKDDATA::iKDDATA / make_shared<KD_TREE> (LidarTypes.cpp:79-83) */
int ml_create(const ml_config *cfg, ml_handle **out);
void ml_destroy(ml_handle *h);
const char *ml_last_error(const ml_handle *h);
/* text for create-time failures (no handle yet) */
const char *ml_global_error(void);

/* ---- map: replaces KD_TREE<PointXYZI> as used through iKDDATA::tree ------------------------------------- */
/* KD_TREE::reset (ikd_Tree.h:262) */
int ml_map_reset(ml_handle *h);
/* KD_TREE::initialized (ikd_Tree.h:263): 1/0 */
int ml_map_initialized(const ml_handle *h);
/* KD_TREE::Build(PointVector) (ikd_Tree.cpp:344-356) <- LidarHelper::init_map_local (LidarHelper.cpp:103-113) */
int ml_map_build(ml_handle *h, const float *xyzi, int n);
/* init_map_local on the current (down-sampled) scan cloud, device-resident (UpdaterLidar.cpp:91-95) */
int ml_map_build_from_scan(ml_handle *h);
/* KD_TREE::Add_Points(PointVector&, bool) with set_downsample_param(ds) (ikd_Tree.cpp:401-479).
 * n_live = live map size afterwards. n_live == NULL (the reference returns nothing): the merge and the kNN-grid rebuild are
 * queued on the handle's stream and the call returns without waiting for the device; the map is current for every later call
 * (stream order), and a capacity overflow - which leaves the map unchanged - is reported as ML_ERR_CAPACITY by the NEXT call on
 * the handle that reads the map (ml_scan_update, ml_map_size, ...). Same rule for ml_map_add_scan and ml_map_reanchor. */
int ml_map_add(ml_handle *h, const float *xyzi, int n, int downsample, float ds, int *n_live);
/* LidarHelper::register_scan (LidarHelper.cpp:196-207): transform the ORIGINAL cloud of the current scan by
 * T_LtoM (set by ml_scan_transform) and Add_Points it. */
int ml_map_add_scan(ml_handle *h, int downsample, float ds, int *n_live);
/* LidarHelper::propagate_map_to_newest_clone (LidarHelper.cpp:209-266) minus the pose look-ups:
 * [VoxelGrid centroid down-sample at ds if do_downsample] -> drop points with intensity + FT + decay_time < now
 * or ||p|| > decay_dist -> p <- T(4x4 fp64, row-major) p -> rebuild. `now_minus_FT` = state->time - FT. */
int ml_map_reanchor(ml_handle *h, const double T_old_to_new[16], double now_minus_FT, double decay_time,
                    double decay_dist, int do_downsample, float ds, int *n_live);
/* KD_TREE::flatten(Root_Node, Storage, NOT_RECORD) (ikd_Tree.cpp:1359-1386): copies min(n, cap) points, *n = live size */
int ml_map_flatten(ml_handle *h, float *xyzi_out, int cap, int *n);
/* KD_TREE::validnum */
int ml_map_size(ml_handle *h, int *n);

/* ---- scan: replaces LiDARData + the LidarHelper pre-processing ------------------------------------------ */
/* LiDARData::LiDARData (LidarTypes.cpp:26-56): range filter (fp32 norm in [rmin, rmax]), intensity = t_rel,
 * fills `pointcloud` and `pointcloud_original`. xyz: n points, `stride` floats apart (3 = packed, 4 = PointXYZ).
 * n_kept may be NULL (then the call does not synchronise). */
int ml_scan_upload(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, int *n_kept);
/* LidarHelper::remove_motion_blur loop body (LidarHelper.cpp:71-85) with per-point transforms interpolated on the
 * device from `n_knots` LiDAR poses T_LitoL (row-major 3x4 fp32 each) sampled uniformly over the sweep:
 * point i uses time fraction (total - i)/total. The reference's inverted guard (:54-58) is NOT reproduced. */
int ml_scan_deskew(ml_handle *h, const float *T_LitoL_knots, int n_knots, int total_points);
/* LidarHelper::downsample (LidarHelper.cpp:89-94): pcl::VoxelGrid<PointXYZI> centroid filter, cubic leaf */
int ml_scan_downsample(ml_handle *h, float leaf, int *n_ds);
/* LidarHelper::transform_to_map tail (LidarHelper.cpp:149-151,192): pointcloud_in_map = T_LtoM * pointcloud (fp32).
 * T row-major 4x4. The transform is remembered for ml_map_add_scan. */
int ml_scan_transform(ml_handle *h, const float T_LtoM[16]);
/* which: 0 = pointcloud, 1 = pointcloud_in_map, 2 = pointcloud_original */
int ml_scan_download(ml_handle *h, int which, float *xyzi_out, int cap, int *n);
/* Storage for queued scans (the role of LiDARData's pcl::PointCloud buffers, LidarTypes.h:37-66): pinned, device-mapped host
 * memory. A scan handed to ml_scan_upload / ml_scan_update from such a buffer (stride 4, 16-byte aligned) is read by the kernels
 * straight over PCIe - no staging memcpy, no DMA copy node; pageable memory works too but pays a host memcpy per scan.
 * Allocation is slow (a driver call): pool the buffers (the C++ shim does). NULL on failure. */
void *ml_host_alloc(size_t bytes);
void ml_host_free(void *p);
/* test hook: replace `pointcloud` (and `pointcloud_original`) verbatim, no filtering */
int ml_scan_set(ml_handle *h, const float *xyzi, int n);

/* ---- ICP pre-alignment: replaces the PointMatcher<float>::ICP block of LidarHelper::transform_to_map -------- */
/* Parameters of the libpointmatcher chain LidarHelper.cpp:156-171 builds (ICP::setDefault, pointmatcher/ICP.cpp:100-113,
 * with the reference filters replaced by MaxDist + SamplingSurfaceNormal). ml_icp_default_options fills the library
 * defaults; the caller sets max_dist = map_icp_dist (OptionsLidar.h:97) and the seed. */
typedef struct ml_icp_options {
  double max_dist;       /* MaxDistDataPointsFilter maxDist: reference = map points with ||p|| < max_dist */
  double read_prob;      /* RandomSamplingDataPointsFilter prob (0.75) */
  double ref_ratio;      /* SamplingSurfaceNormalDataPointsFilter ratio (0.5) */
  int normal_knn;        /* SamplingSurfaceNormalDataPointsFilter knn (7): support of the PCA normal */
  double trim_ratio;     /* TrimmedDistOutlierFilter ratio (0.85) */
  int max_iter;          /* CounterTransformationChecker maxIterationCount (40) */
  double min_diff_rot;   /* DifferentialTransformationChecker minDiffRotErr (0.001 rad) */
  double min_diff_trans; /* DifferentialTransformationChecker minDiffTransErr (0.001 m) */
  int smooth_len;        /* DifferentialTransformationChecker smoothLength (3) */
  double max_match_dist; /* bound of the 1-NN matcher [m]. The reference's KDTreeMatcher is unbounded (maxDist = inf) and
                            leaves far matches to the trimmed filter; a bound keeps the grid walk local. Reading points
                            without a reference point within the bound count as unmatched (weight 0). Default 2. */
  double normal_radius;  /* neighbours for the normal are searched within this radius [m]; default 2 */
  uint64_t seed;         /* the reference seeds its samplers from std::random_device / std::rand (non-deterministic) */
} ml_icp_options;
typedef struct ml_icp_result {
  float T_LtoM[16];      /* what transform_to_map stores in lidar->T_LtoM: T_icp on success, else the initial guess */
  float T_icp[16];       /* raw ICP output */
  int success;           /* lidar->icp_success: !maxNumIterationsReached && ||trans(T_icp * T_init^-1)|| <= 1 (LidarHelper.cpp:177) */
  int iterations;
  int max_iter_reached;
  int n_reading, n_reference; /* points after the reading / reference filters */
  int n_matched, n_used;      /* last iteration: matches within the bound, matches with weight 1 */
  double mean_sq_err;         /* last iteration: mean squared point-to-plane error over the used matches, before its update */
} ml_icp_result;
void ml_icp_default_options(ml_icp_options *o);
/* PointMatcher<float>::ICP::compute(new_points, map_points, T_LtoM) (LidarHelper.cpp:173-184) on the current scan cloud
 * (`pointcloud`, after ml_scan_downsample) against the current map. The caller then passes res->T_LtoM to
 * ml_scan_transform, like LidarHelper.cpp:192. Returns ML_SOFT_NO_ROWS when either cloud is (nearly) empty. */
int ml_scan_icp(ml_handle *h, const float T_init[16], const ml_icp_options *o, ml_icp_result *res);

/* ---- simulator LiDAR model: replaces Simulator::get_lidar_pointcloud ------------------- */
/* The world of Simulator::load_plane_data (mins/src/sim/Simulator.cpp:990-1096): n_planes x 22 doubles
 * [A B C D | top_left(3) | bottom_right(3) | V1 V2 V3 V4 (3 each)] with the plane equation and the edge directions of
 * SimulationPlane (mins/src/sim/SimulationPlane.h:43-53,80-87). */
int ml_sim_set_world(ml_handle *h, const double *planes, int n_planes);
/* Ray bearings in the LiDAR frame, n_rays x 3 doubles, in the reference's emission order (for v_angle: for h_angle,
 * Simulator.cpp:756-757,767-769). n_rays <= max_scan_points. */
int ml_sim_set_rays(ml_handle *h, const double *dirs_L, int n_rays);
/* Simulator::get_lidar_pointcloud (Simulator.cpp:742-815) for one sensor pose: p_LinG = p_IinG + R_GtoI^T p_LinI (:765),
 * R_LtoG = R_GtoI^T R_ItoL^T (:769, row-major). Nearest hit over all planes (SimulationPlane.h:73-95), rays outside
 * [min_range, max_range] dropped (:788), p = d * bearing cast to fp32 (:792-795), + noise * N(0,1) per axis in fp32
 * (:798-803; noise <= 0: none). The reference draws from a sequential std::mt19937; here the normal deviates come from
 * Philox-4x32-10 keyed by `seed` with the ray index as counter (order-independent). The scan stays on the device (it
 * becomes the raw-scan buffer, see ml_sim_scan_device); xyz_out (stride 4, may be NULL) receives min(cap, n) points. */
int ml_sim_raycast(ml_handle *h, const double p_LinG[3], const double R_LtoG[9], double min_range, double max_range, float noise,
                   uint64_t seed, float *xyz_out, int cap, int *n_out);
/* The same with the motion distortion of Simulator.cpp:761-771 (`raw_remove_motion_blur`): ray i of n_rays is cast from the
 * sensor pose at time - (n_rays - i) * raw_point_dt. pose_knots: n_knots >= 2 poses [R_LtoG row-major (9) | p_LinG (3)] sampled
 * uniformly over [time - n_rays * raw_point_dt, time]; a ray blends its two bracketing knots linearly. The inverse operation
 * on the estimator side is ml_scan_deskew. */
int ml_sim_raycast_sweep(ml_handle *h, const double *pose_knots, int n_knots, double min_range, double max_range, float noise,
                         uint64_t seed, float *xyz_out, int cap, int *n_out);
/* device pointer (stride 4) and size of the last ray-cast scan: pass them to ml_scan_upload / ml_scan_update, which use
 * device-resident input in place. Valid until the next upload of a host scan. */
int ml_sim_scan_device(ml_handle *h, const float **xyz_dev, int *n);

/* ---- update: replaces the point loop + stacking + compression of UpdaterLidar::update ------------------- */
/* LidarHelper::get_neighbors / KD_TREE::Nearest_Search (LidarHelper.cpp:268-289, ikd_Tree.cpp:358-389) for every
 * point of pointcloud_in_map. Exact k-NN on fp32 squared distance ((dx*dx + dy*dy) + dz*dz, no FMA), total order
 * (d2, map index) ascending. idx: n*k int32 (index into ml_map_flatten order; -1 where fewer than k were found
 * within sqrt(max_d2)), d2: n*k float. ok[i] = 1 iff get_neighbors would return true. Any output may be NULL. */
int ml_knn(ml_handle *h, int k, float max_d2, int32_t *idx, float *d2, uint8_t *ok, int *n);
/* UpdaterLidar.cpp:269-397 for the whole scan: match -> plane -> residual/Jacobian -> whitening -> nullspace
 * projection -> chi2 gate -> reduction. Returns the information form of the stacked, projected system:
 * HtH = H_big^T H_big (total_hx x total_hx, row-major), Htr = H_big^T res_big (total_hx), R = I.
 * P: marginal covariance of H_order (total_hx x total_hx, row-major; UpdaterLidar.cpp:257).
 * Returns ML_SOFT_NO_ROWS when no point was accepted. */
int ml_linsys(ml_handle *h, const ml_scan_consts *c, const double *P, double *HtH, double *Htr, ml_counters *cnt);
/* One call for the timed path: ml_scan_upload + ml_scan_transform + ml_linsys with a single synchronisation. */
int ml_scan_update(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax,
                   float leaf /* <=0: no down-sampling */, const float T_LtoM[16], const ml_scan_consts *c,
                   const double *P, double *HtH, double *Htr, ml_counters *cnt);
/* parity export of the last ml_linsys: per-point status/plane/residual/chi2 and the whitened Jacobian row h
 * (n x total_hx, row-major; may be NULL). */
int ml_point_rows(ml_handle *h, ml_point_row *rows, double *h_rows, int cap, int *n);

/* host-side tail (plain C++, no device work): turn (HtH, Htr) into the compressed (R_c, r_c) that
 * StateHelper::measurement_compress_inplace would hand to EKFUpdate (StateHelper.cpp:631-643): R_c^T R_c = HtH,
 * R_c^T r_c = Htr, rows with zero pivot dropped. Rc: total_hx x total_hx row-major, rc: total_hx. Returns rank. */
int ml_info_to_sqrt(const double *HtH, const double *Htr, int total_hx, double *Rc, double *rc);

/* device-time (ms, CUDA events on the handle's stream) of the stages of the last ml_linsys / ml_scan_update:
 * [0] upload+range filter (+ transform when not down-sampling), [1] downsample, [2] transform, [3] query binning, [4] knn,
 * [5] rows + reduction + finalize (one fused kernel), [6] 0 (the kernel of [5] writes the result block straight into pinned
 * host memory; there is no separate D2H copy), [7] total */
int ml_last_timing(ml_handle *h, float ms[8]);
/* CUDA-event timers: 0 = off, 1 = per-stage (default; diagnostic - the event records between kernels serialise the
 * programmatically dependent launches of the pipeline), 2 = total only ([7]; the other entries read 0) */
int ml_set_timing(ml_handle *h, int on);
/* profiling aid, not part of the reference's surface: per-CTA %globaltimer stamps [ns] of the kNN and rows kernels of the scans
 * that follow (on = 1 arms, 0 disarms); out (may be NULL) receives the stamps of the last scan as [kernel: 0 kNN, 1 rows][cta <
 * 2048][slot < 8] and clears them. Returns the number of values written. Costs one predicated store per mark when disarmed. */
int ml_debug_timeline(ml_handle *h, int on, unsigned long long *out, int cap);
/* Residual statistics of UpdaterStatistics::Chi2Check (res_stat->add_stat(res), UpdaterStatistics.cpp:54-57): on = 1 makes
 * ml_linsys / ml_scan_update also return ml_counters.sum_res - the first moment of the projected residuals in the reference's own
 * nullspace basis (StateHelper::Givens_Rotation, StateHelper.cpp:660-680), run per point on the device. Default off. */
int ml_set_stats(ml_handle *h, int on);
/* ---- allocation-free range filter -------------------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it). The range filter of ml_scan_upload (LiDARData's
 * constructor, fp32 norm in [rmin, rmax], intensity = t_rel, order kept) without a handle: the call allocates nothing and frees
 * nothing; its only scratch (one count per 2048-point tile) is a static device array of ML_STATIC_MAX_POINTS / 2048 words (8 KB)
 * that every GPU loading the library holds until the process exits. n > ML_STATIC_MAX_POINTS returns ML_ERR_CAPACITY.
 * xyz_dev: n points, `stride` floats apart (3 or 4), memory the CURRENT device can read; xyzi_out_dev: room for n points of 4
 * floats, 16-byte aligned, device memory; n_kept_dev: one int in device memory. Two kernels are queued on `stream` (a
 * cudaStream_t, NULL = the default stream) and the call returns without synchronising. Every call on a device shares the scratch
 * array, which keeps the previous call's contents (nothing in it needs clearing: a call overwrites what it reads): calls that
 * overlap on the GPU - different streams, concurrent graph launches - race. The first call on a device looks the array's
 * address up (cudaGetSymbolAddress): make it before a stream capture begins. Errors: ml_global_error(). */
#define ML_STATIC_MAX_POINTS (1 << 22)
int ml_range_filter_static(const float *xyz_dev, int stride, int n, float t_rel, float rmin, float rmax, float *xyzi_out_dev,
                           int *n_kept_dev, void *stream);

/* sizeof of {ml_scan_consts, ml_point_row, ml_counters, ml_config} as compiled, for FFI layout checks */
void ml_abi_sizes(int out[4]);
/* number of kernel launches issued by this handle so far */
long long ml_launch_count(const ml_handle *h);

#ifdef __cplusplus
}
#endif
#endif
