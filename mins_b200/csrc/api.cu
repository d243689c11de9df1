// C ABI of the B200-native MINS LiDAR update path (see include/mins_lidar.h). Host-side orchestration only:
// every numeric step is a hand-written sm_100a kernel in map.cuh / scan.cuh / update.cuh / rows.cuh.
// There is no CPU fallback: without a usable CUDA device ml_create fails.
#include "common.cuh"
#include "map.cuh"
#include "prims.cuh"
#include "rows.cuh"
#include "scan.cuh"
#include "sim.cuh"
#include "update.cuh"
#include "fused.cuh"
#include "tile.cuh"
#include "icp.cuh"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <utility>
#include <vector>

using namespace ml;

static std::string g_global_error;
#if defined(ML_KNN_PROF) && ML_KNN_PROF
#define ML_TL_MAX_CTAS 8192
#else
#define ML_TL_MAX_CTAS 2048
#endif

struct ml_handle {
  ml_config cfg;
  cudaStream_t stream = nullptr;
  int n_sm = 148;
  int knn_calls = 0; // parity of the kNN work-list counters (ScanState::n_units)
  std::string err;
  long long launches = 0;
  int cap_scan = 0, cap_map = 0, cap_cells = 0, cap_all = 0;
  double grid_cell = 0.5;
  // map
  float4 *map_pts[2] = {nullptr, nullptr};
  int cur = 0;
  int *d_map_n = nullptr;
  int map_n = 0;
  bool map_init = false;
  GridDesc *d_grid = nullptr;
  uint32_t *d_bbox = nullptr;
  uint32_t *d_bbox_vg = nullptr; // the voxel filter's own box: in its initial (empty) state between uses
  uint32_t *cell_cnt = nullptr, *cell_start = nullptr;
  float4 *cell_pts = nullptr;
  uint32_t *pt_cell = nullptr, *pt_rank = nullptr;
  uint32_t *tile_sums = nullptr; // per-tile counts of the two-launch compactions (range filter fallback, simulator)
  VoxSlot *vox_tab = nullptr;
  int vox_slots = 0;
  uint32_t *add_slot = nullptr, *contender = nullptr, *keep = nullptr, *pos = nullptr;
  int *d_ints = nullptr; // scratch ints: [0]=n_add [1]=total [2]=overflow [3]=n_tmp [4]=passthrough
  float4 *add_pts = nullptr;
  // scan
  float *d_raw = nullptr;
  float *h_raw = nullptr; // pinned staging (device-mapped: h_raw_dev)
  const float *h_raw_dev = nullptr;
  float4 *scan_pts = nullptr, *scan_orig = nullptr, *scan_map = nullptr, *scan_tmp = nullptr;
  ScanState *d_st = nullptr;
  // decoupled look-back (prims.cuh), shared by every single-pass scan/compaction on the handle's stream: tile states
  // (epoch-tagged, never reset), tile ticket (back at 0 after each launch), epoch of the last launch
  unsigned long long *lb_state = nullptr;
  unsigned int *lb_ctr = nullptr;
  unsigned int lb_epoch = 0;
  unsigned int *mg_done = nullptr;  // k_map_merge: CTAs finished
  MergeResult *h_merge = nullptr;   // pinned: result block of the last k_map_merge
  bool merge_pending = false;       // a k_map_merge is in flight: map_n / overflow not yet read back (resolve_merge)
  const char *merge_who = "";
  bool bin_done = false; // the queries of the resident scan are already in the hash table (fused front end)
  bool vg_ready = false; // d_vg already holds the voxel grid of the resident scan for the leaf of this call (fused front end)
  int n_raw = 0;     // host-known upper bound of the scan size
  bool have_T = false;
  Tf32 T_LtoM;
  uint32_t *sort_keys[2] = {nullptr, nullptr}, *sort_vals[2] = {nullptr, nullptr}, *radix_ghist = nullptr, *radix_dbase = nullptr;
  unsigned long long *radix_state = nullptr; // per (tile, digit) look-back states of k_radix_pass, epoch-tagged
  int radix_tiles_cap = 0;
  VoxGridDesc *d_vg = nullptr;
  float *d_knots = nullptr;
  size_t knots_cap = 0;
  // simulator (ml_sim_*)
  double *sim_planes = nullptr, *sim_dirs = nullptr, *sim_dist = nullptr;
  int sim_n_planes = 0, sim_n_rays = 0;
  int *sim_n_out = nullptr;
  int sim_last_n = 0;
  double *sim_knots = nullptr;
  size_t sim_knots_cap = 0;
  // ICP pre-alignment (ml_scan_icp), allocated on first use
  uint8_t *icp_keep = nullptr;
  float4 *icp_normals = nullptr, *icp_cell_pts = nullptr, *icp_read = nullptr, *icp_q = nullptr;
  IcpState *d_icp = nullptr;
  ScanState *d_ist = nullptr;
  double *icp_parts = nullptr;
  IcpHostOut *h_icp = nullptr;
  int icp_knn_calls = 0;
  // update
  QSlot *qtab = nullptr;
  int q_slots = 0;
  uint32_t *q_slot = nullptr, *q_rank = nullptr;
  float4 *q_sorted = nullptr;
  KnnUnit *units = nullptr;
  int units_cap = 0;
  int32_t *knn_idx = nullptr;
  float *knn_d2 = nullptr;
  uint8_t *knn_ok = nullptr;
  float4 *nbr = nullptr; // k neighbour coordinates per query, scan order (written by k_knn, read by k_rows_reduce)
  RowConsts rowc;        // per-scan constants of the rows kernel (kernel parameter), filled by consts_async
  double *d_T = nullptr; // T on the device: parity export only (k_expand_h)
  PointRow *rows = nullptr;
  RRPart *h_parts = nullptr; // pinned, device-mapped: one reduced system per CTA of the rows kernel (rows.cuh)
  int rows_nb = 0;           // CTAs of the last rows launch
  // fused front end + kNN (fused.cuh)
  FusedCtl *d_fctl = nullptr;
  uint32_t *units_tag = nullptr;
  bool use_fused_knn = true; // MINS_FUSED_KNN=0: the separate kernels (A/B)
  // the one-kernel path (tile.cuh)
  bool use_tile = true;      // MINS_SCAN_TILE=0: front + kNN kernel followed by the rows kernel (A/B); 2: always (see ml_scan_update)
  bool tile_force = false;
  bool tile_keep_nbr = false; // MINS_TILE_NBR=1: k_scan_tile also leaves the neighbours' coordinates in `nbr` (10 MB of stores per
                              // 131 072-ray scan, 2.6 us); by default ml_point_rows regenerates them with the multi-call search
  bool rows_need_knn = false; // the resident scan's rows are exportable but `nbr` has to be produced first (ml_point_rows)
  bool tile_poll = true;     // MINS_TILE_POLL=0: cudaStreamSynchronize before the parts are read, instead of watching their words land
  int tile_grid[ML_MAX_K + 1][2][2] = {}; // co-resident CTAs per (k, NB, VEC4), 0 = attribute not set yet
  int parts_armed = 0, parts_armed_nred = 0; // parts [0, parts_armed) hold the sentinel in their first parts_armed_nred doubles and counters
  int parts_n = 0;           // parts of the last reduction (rows_host_tail)
  bool host_prof = false; // MINS_HOST_PROF=1
  double host_prof_us[4] = {0, 0, 0, 0};
  long host_prof_n = 0;
  int fused_producers = 0;   // MINS_FUSED_PRODUCERS: 0 = by input kind
  bool fused_attr[ML_MAX_K + 1][2] = {};
  bool rows_exported = false;
  bool rows_attr[ML_MAX_K + 1][8] = {};
  // debug timeline (ml_debug_timeline): per-CTA %globaltimer stamps of the kNN and rows kernels
  bool stats_on = false; // ml_set_stats
  bool tl_on = false;
  unsigned long long *tl_rows = nullptr, *tl_knn = nullptr;
  int rows_block = 0; // 0 = chosen per scan size; MINS_ROWS_BLOCK=512|1024 pins it (profiling) // cudaFuncAttributeMaxDynamicSharedMemorySize set for k_rows_reduce<K, NB> on THIS handle's device
  double *d_hrows = nullptr;            // parity export (ml_point_rows): n x total_hx, grown on demand, never on the per-scan path
  size_t hrows_cap = 0;
  // per-scan constants, host side: the caller's struct, T (h = T^T g) and the result of the last rows launch
  ml_scan_consts consts;
  double hT[ML_G * ML_MAX_HX];
  int last_n_scan = 0;
  bool rows_valid = false;
  int rows_k = 0;
  // timing
  cudaEvent_t ev[10];
  int timing = 1; // 0 off, 1 per-stage events (diagnostic: event records between kernels defeat the launch overlap), 2 total only
  bool use_pdl = true;
  bool use_fused_front = true;
  bool zero_copy = true; // MINS_ZERO_COPY=0 restores the DMA copy of pinned input
  float last_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

#define CK(call)                                                                                                       \
  do {                                                                                                                 \
    cudaError_t e__ = (call);                                                                                          \
    if (e__ != cudaSuccess) {                                                                                          \
      char buf__[512];                                                                                                 \
      snprintf(buf__, sizeof buf__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__);      \
      if (h)                                                                                                           \
        h->err = buf__;                                                                                                \
      else                                                                                                             \
        g_global_error = buf__;                                                                                        \
      return ML_ERR_CUDA;                                                                                              \
    }                                                                                                                  \
  } while (0)

#define LAUNCH(kern, grid, block, ...)                                                                                 \
  do {                                                                                                                 \
    kern<<<(grid), (block), 0, h->stream>>>(__VA_ARGS__);                                                              \
    h->launches++;                                                                                                     \
  } while (0)

// launch on the handle's stream with programmatic stream serialization (see pdl_enter in common.cuh): the kernel may be
// scheduled while its predecessor drains; it calls griddepcontrol.wait before touching memory.
template <typename... KArgs, typename... Args>
static inline void launch_pdl(ml_handle *h, void (*kern)(KArgs...), int grid, int block, size_t smem, Args &&...args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid), cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = h->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = h->use_pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
  h->launches++;
}
#define LAUNCHP(kern, grid, block, ...) launch_pdl(h, kern, (grid), (block), 0, __VA_ARGS__)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline int *st_field(ml_handle *h, size_t off) { return (int *)((char *)h->d_st + off); }
#define ST(f) st_field(h, offsetof(ScanState, f))

static int fail(ml_handle *h, int code, const char *msg) {
  if (h)
    h->err = msg;
  else
    g_global_error = msg;
  return code;
}

// the fused front end leaves the scan's queries in the hash table for the kNN that normally follows; if that kNN never ran
// (error path, new upload), the table must be emptied before anybody bins into it again
static void drop_stale_bins(ml_handle *h) {
  if (h->bin_done) {
    cudaMemsetAsync(h->qtab, 0, sizeof(QSlot) * (size_t)h->q_slots, h->stream);
    h->bin_done = false;
  }
}

static int check_launch(ml_handle *h) {
  CK(cudaGetLastError());
  return ML_OK;
}

static uint32_t lb_next_epoch(ml_handle *h) {
  h->lb_epoch = (h->lb_epoch + 1u) & 0x3fffffffu;
  if (h->lb_epoch == 0u)
    h->lb_epoch = 1u;
  return h->lb_epoch;
}

// exclusive scan of `in` (n on device) into out[0..n], one pass; total optionally to d_total. n_ub = host upper bound.
static void excl_scan(ml_handle *h, uint32_t *in, const int *d_n, int n_ub, uint32_t *out, int *d_total) {
  if (n_ub > (1 << 18)) // long inputs want large tiles (scripts/bench_scan.cu: 1.1 M elements 12.6 us vs 10.5 us)
    LAUNCH((k_scan_lb<SCAN_ITEMS_LONG, false>), cdiv(n_ub, SCAN_TILE_LONG), SCAN_BLOCK, in, d_n, out, d_total, h->lb_state, h->lb_ctr,
           lb_next_epoch(h));
  else
    LAUNCH((k_scan_lb<SCAN_ITEMS, false>), std::max(1, cdiv(n_ub, SCAN_TILE)), SCAN_BLOCK, in, d_n, out, d_total, h->lb_state, h->lb_ctr,
           lb_next_epoch(h));
}

// stable sort of (keys, vals) by key; result ends in sort_keys[0]/sort_vals[0]. bits = number of key bits to sort.
// stable sort of (sort_keys[0], sort_vals[0]) by key, `bits` key bits; the result ends in buffer 0. The digit bases
// (radix_dbase) must have been produced with the keys (k_vox_index_hist / k_radix_hist4).
static void radix_sort(ml_handle *h, const int *d_n, int n_ub, int bits) {
  int passes = std::min(RS_PASSES, std::max(1, (bits + 7) / 8));
  if (passes & 1)
    passes++; // even number of passes so the result lands in buffer 0
  const int tiles = std::max(1, cdiv(n_ub, RS_TILE));
  int src = 0;
  for (int p = 0; p < passes; p++) {
    LAUNCH(k_radix_pass<RS_ITEMS>, tiles, RS_BLOCK, h->sort_keys[src], h->sort_vals[src], d_n, p, h->radix_dbase, h->radix_state, h->lb_ctr,
           lb_next_epoch(h), h->sort_keys[1 - src], h->sort_vals[1 - src]);
    src = 1 - src;
  }
}

// dense kNN grid over map_pts[cur] from the descriptor in d_grid: cell histogram -> one-pass scan over the cells -> scatter.
// Nothing here waits for the device: launches are sized by host-side upper bounds (n_ub points, cap_cells cells), the device
// descriptor is authoritative and surplus CTAs leave at once.
static int grid_build_tail(ml_handle *h, int n_ub) {
  if (n_ub <= 0)
    return ML_OK;
  LAUNCH(k_cell_count, cdiv(n_ub, 256), 256, h->map_pts[h->cur], h->d_grid, h->cell_cnt, h->pt_cell, h->pt_rank);
  // cell_start = exclusive scan of the counters, which the same pass leaves all-zero for the next build
  LAUNCH((k_scan_lb<SCAN_ITEMS_LONG, true>), std::max(1, cdiv(h->cap_cells, SCAN_TILE_LONG)), SCAN_BLOCK, h->cell_cnt, &h->d_grid->ncells,
         h->cell_start, (int *)nullptr, h->lb_state, h->lb_ctr, lb_next_epoch(h));
  LAUNCH(k_cell_scatter, cdiv(n_ub, 256), 256, h->map_pts[h->cur], h->d_grid, h->cell_start, h->pt_cell, h->pt_rank, h->cell_pts);
  return check_launch(h);
}

// (re)build the dense kNN grid over map_pts[cur] (n = *d_map_n, host copy h->map_n)
static int rebuild_grid(ml_handle *h) {
  const int n = h->map_n;
  LAUNCH(k_bbox_init, 1, 32, h->d_bbox);
  if (n > 0)
    LAUNCH(k_bbox, std::min(148 * 8, cdiv(n, 256)), 256, h->map_pts[h->cur], h->d_map_n, h->d_bbox);
  LAUNCH(k_grid_setup, 1, 32, h->d_bbox, h->d_map_n, h->grid_cell, h->cap_cells, h->d_grid);
  h->map_init = n > 0;
  return grid_build_tail(h, n);
}

// Waits for the merge in flight (if any) and takes over its result block: the new point count, or - when the capacity was
// exceeded - the previous map (double-buffered, untouched) with its grid rebuilt and ML_ERR_CAPACITY. Every entry point
// that reads the host-side map state calls this first; a caller that did not ask for the live count (n_live == NULL) thus
// gets its map update queued behind the scan's kernels and overlapped with its own host work.
static int resolve_merge(ml_handle *h) {
  if (!h->merge_pending)
    return ML_OK;
  h->merge_pending = false;
  CK(cudaStreamSynchronize(h->stream));
  if (h->h_merge->overflow) {
    h->cur = 1 - h->cur;
    int rc = rebuild_grid(h);
    h->err = std::string(h->merge_who) + ": map capacity (max_map_points) exceeded; the map is unchanged";
    return rc ? rc : ML_ERR_CAPACITY;
  }
  h->map_n = h->h_merge->n_out;
  h->map_init = h->map_n > 0;
  return ML_OK;
}
// the calling thread may have another device current (two LiDARs on two GPUs in one process): every entry point that
// touches the device makes the handle's device current first
static inline void use_device(const ml_handle *h) {
  int d = -1;
  if (cudaGetDevice(&d) != cudaSuccess || d != h->cfg.device)
    cudaSetDevice(h->cfg.device);
}
#define RESOLVE(h)                                                                                                      \
  do {                                                                                                                 \
    use_device(h);                                                                                                     \
    if (int rc_ = resolve_merge(h))                                                                                    \
      return rc_;                                                                                                      \
  } while (0)

// one-pass merge of the survivors of [a | b] into the other map buffer + bbox + grid descriptor (k_map_merge), then the grid
// over it. Flips the double buffer at once (stream order makes every later launch see the new map); the host-side count and
// the capacity verdict arrive with resolve_merge: now if `wait`, else at the next entry point that needs them.
template <int MODE>
static int merge_into_map(ml_handle *h, const float4 *a, const int *d_na, const float4 *b, const int *d_nb, int n_ub, const char *who,
                          bool wait) {
  LAUNCH(k_bbox_init, 1, 32, h->d_bbox);
  LAUNCH(k_map_merge<MODE>, std::max(1, cdiv(n_ub, MG_TILE)), MG_BLOCK, a, d_na, b, d_nb, h->vox_tab, h->contender, h->add_slot, h->keep,
         h->map_pts[1 - h->cur], h->cap_map, h->lb_state, h->lb_ctr, lb_next_epoch(h), h->d_bbox, h->mg_done, h->d_ints + 3, h->grid_cell,
         h->cap_cells, h->d_grid, h->d_map_n, h->h_merge);
  h->cur = 1 - h->cur;
  int rc = grid_build_tail(h, std::min(n_ub, h->cap_map));
  h->merge_pending = true;
  h->merge_who = who;
  if (rc || wait) {
    int rc2 = resolve_merge(h);
    return rc ? rc : rc2;
  }
  return ML_OK;
}

extern "C" {

const char *ml_global_error(void) { return g_global_error.c_str(); }
const char *ml_last_error(const ml_handle *h) { return h ? h->err.c_str() : g_global_error.c_str(); }
long long ml_launch_count(const ml_handle *h) { return h ? h->launches : 0; }

int ml_create(const ml_config *cfg, ml_handle **out) {
  ml_handle *h = nullptr;
  if (!cfg || !out || cfg->max_scan_points <= 0 || cfg->max_map_points <= 0)
    return fail(nullptr, ML_ERR_ARG, "ml_create: bad config");
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (ndev <= 0 || cfg->device < 0 || cfg->device >= ndev)
    return fail(nullptr, ML_ERR_CUDA, "ml_create: no usable CUDA device (this library has no CPU path)");
  CK(cudaSetDevice(cfg->device));
  h = new ml_handle();
  h->cfg = *cfg;
  h->cap_scan = cfg->max_scan_points;
  h->cap_map = cfg->max_map_points;
  h->cap_cells = cfg->max_grid_cells > 0 ? cfg->max_grid_cells : (1 << 24);
  h->cap_all = h->cap_map + h->cap_scan;
  h->grid_cell = cfg->grid_cell > 0 ? (double)cfg->grid_cell : 0.5;
  if (const char *e = getenv("MINS_NO_PDL")) // A/B switches for profiling
    h->use_pdl = !(e[0] == '1');
  if (const char *e = getenv("MINS_ZERO_COPY"))
    h->zero_copy = !(e[0] == '0');
  if (const char *e = getenv("MINS_ROWS_BLOCK"))
    h->rows_block = atoi(e);
  if (const char *e = getenv("MINS_FUSED_KNN"))
    h->use_fused_knn = !(e[0] == '0');
  if (const char *e = getenv("MINS_FUSED_PRODUCERS"))
    h->fused_producers = atoi(e);
  if (const char *e = getenv("MINS_SCAN_TILE")) {
    h->use_tile = !(e[0] == '0');
    h->tile_force = e[0] == '2';
  }
  if (const char *e = getenv("MINS_TILE_NBR"))
    h->tile_keep_nbr = !(e[0] == '0');
  if (const char *e = getenv("MINS_TILE_POLL"))
    h->tile_poll = !(e[0] == '0');
  if (const char *e = getenv("MINS_HOST_PROF"))
    h->host_prof = e[0] == '1';
  if (const char *e = getenv("MINS_NO_FUSED_FRONT"))
    h->use_fused_front = !(e[0] == '1');
  CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  CK(cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, cfg->device));
  if (h->n_sm > ML_NPART_MAX)
    h->n_sm = ML_NPART_MAX;
  for (auto &e : h->ev)
    CK(cudaEventCreate(&e));
  const size_t cs = (size_t)h->cap_scan, ca = (size_t)h->cap_all;
  CK(cudaMalloc(&h->map_pts[0], sizeof(float4) * ca));
  CK(cudaMalloc(&h->map_pts[1], sizeof(float4) * ca));
  CK(cudaMalloc(&h->d_map_n, sizeof(int)));
  CK(cudaMemset(h->d_map_n, 0, sizeof(int)));
  CK(cudaMalloc(&h->d_grid, sizeof(GridDesc)));
  CK(cudaMemset(h->d_grid, 0, sizeof(GridDesc)));
  CK(cudaMalloc(&h->d_bbox, sizeof(uint32_t) * 8));
  CK(cudaMalloc(&h->d_bbox_vg, sizeof(uint32_t) * 8));
  k_bbox_init<<<1, 32, 0, h->stream>>>(h->d_bbox_vg);
  CK(cudaMalloc(&h->cell_cnt, sizeof(uint32_t) * ((size_t)h->cap_cells + 2)));
  CK(cudaMalloc(&h->cell_start, sizeof(uint32_t) * ((size_t)h->cap_cells + 2)));
  CK(cudaMalloc(&h->cell_pts, sizeof(float4) * ca));
  CK(cudaMalloc(&h->pt_cell, sizeof(uint32_t) * ca));
  CK(cudaMalloc(&h->pt_rank, sizeof(uint32_t) * ca));
  CK(cudaMemset(h->cell_cnt, 0, sizeof(uint32_t) * ((size_t)h->cap_cells + 2))); // all-zero between grid builds (k_scan_lb<.., true>)
  CK(cudaMalloc(&h->tile_sums, sizeof(uint32_t) * (cs / RF_TILE + 4)));
  h->vox_slots = 1;
  while ((size_t)h->vox_slots < 2 * cs)
    h->vox_slots <<= 1;
  CK(cudaMalloc(&h->vox_tab, sizeof(VoxSlot) * (size_t)h->vox_slots));
  CK(cudaMalloc(&h->add_slot, sizeof(uint32_t) * cs));
  CK(cudaMalloc(&h->contender, sizeof(uint32_t) * ca));
  CK(cudaMalloc(&h->keep, sizeof(uint32_t) * (ca + 2)));
  CK(cudaMalloc(&h->pos, sizeof(uint32_t) * (ca + 2)));
  CK(cudaMalloc(&h->d_ints, sizeof(int) * 16));
  CK(cudaMemset(h->d_ints, 0, sizeof(int) * 16));
  CK(cudaMalloc(&h->add_pts, sizeof(float4) * cs));
  CK(cudaMalloc(&h->d_raw, sizeof(float) * 4 * cs));
  CK(cudaHostAlloc(&h->h_raw, sizeof(float) * 4 * cs, cudaHostAllocMapped));
  {
    void *dp = nullptr;
    if (cudaHostGetDevicePointer(&dp, h->h_raw, 0) == cudaSuccess)
      h->h_raw_dev = (const float *)dp;
    else
      cudaGetLastError();
  }
  CK(cudaMalloc(&h->scan_pts, sizeof(float4) * cs));
  CK(cudaMalloc(&h->scan_orig, sizeof(float4) * cs));
  CK(cudaMalloc(&h->scan_map, sizeof(float4) * cs));
  CK(cudaMalloc(&h->scan_tmp, sizeof(float4) * ca));
  CK(cudaMalloc(&h->d_st, sizeof(ScanState)));
  CK(cudaMemset(h->d_st, 0, sizeof(ScanState)));
  const size_t lb_tiles = std::max(std::max((size_t)h->cap_cells / SCAN_TILE, ca / SCAN_TILE), std::max(ca / MG_TILE, cs / RF_TILE)) + 4;
  CK(cudaMalloc(&h->lb_state, sizeof(unsigned long long) * lb_tiles));
  CK(cudaMemset(h->lb_state, 0, sizeof(unsigned long long) * lb_tiles));
  CK(cudaMalloc(&h->lb_ctr, sizeof(unsigned int)));
  CK(cudaMemset(h->lb_ctr, 0, sizeof(unsigned int)));
  CK(cudaMalloc(&h->mg_done, sizeof(unsigned int)));
  CK(cudaMemset(h->mg_done, 0, sizeof(unsigned int)));
  CK(cudaMallocHost(&h->h_merge, sizeof(MergeResult)));
  memset(h->h_merge, 0, sizeof(MergeResult));
  for (int b = 0; b < 2; b++) {
    CK(cudaMalloc(&h->sort_keys[b], sizeof(uint32_t) * ca));
    CK(cudaMalloc(&h->sort_vals[b], sizeof(uint32_t) * ca));
  }
  h->radix_tiles_cap = cdiv((long long)ca, RS_TILE) + 1;
  CK(cudaMalloc(&h->radix_state, sizeof(unsigned long long) * 256 * (size_t)h->radix_tiles_cap));
  CK(cudaMemset(h->radix_state, 0, sizeof(unsigned long long) * 256 * (size_t)h->radix_tiles_cap));
  CK(cudaMalloc(&h->radix_ghist, sizeof(uint32_t) * 256 * RS_PASSES));
  CK(cudaMemset(h->radix_ghist, 0, sizeof(uint32_t) * 256 * RS_PASSES));
  CK(cudaMalloc(&h->radix_dbase, sizeof(uint32_t) * 256 * RS_PASSES));
  CK(cudaMalloc(&h->d_vg, sizeof(VoxGridDesc)));
  h->q_slots = 1;
  while ((size_t)h->q_slots < 2 * cs)
    h->q_slots <<= 1;
  CK(cudaMalloc(&h->qtab, sizeof(QSlot) * (size_t)h->q_slots));
  CK(cudaMemset(h->qtab, 0, sizeof(QSlot) * (size_t)h->q_slots)); // kept clean by k_knn from then on
  CK(cudaMalloc(&h->q_slot, sizeof(uint32_t) * cs));
  CK(cudaMalloc(&h->q_rank, sizeof(uint32_t) * cs));
  CK(cudaMalloc(&h->q_sorted, sizeof(float4) * cs));
  h->units_cap = (int)(cs + cs / 32 + 64);
  CK(cudaMalloc(&h->units, sizeof(KnnUnit) * (size_t)h->units_cap));
  CK(cudaMalloc(&h->knn_idx, sizeof(int32_t) * cs * ML_MAX_K));
  CK(cudaMalloc(&h->knn_d2, sizeof(float) * cs * ML_MAX_K));
  CK(cudaMalloc(&h->knn_ok, cs));
  CK(cudaMalloc(&h->nbr, sizeof(float4) * cs * ML_MAX_K));
  CK(cudaMalloc(&h->d_T, sizeof(double) * ML_G * ML_MAX_HX));
  CK(cudaMalloc(&h->d_fctl, sizeof(FusedCtl)));
  CK(cudaMemset(h->d_fctl, 0, sizeof(FusedCtl)));
  CK(cudaMalloc(&h->units_tag, sizeof(uint32_t) * ((size_t)h->units_cap + 16384)));
  CK(cudaMemset(h->units_tag, 0, sizeof(uint32_t) * ((size_t)h->units_cap + 16384)));
  CK(cudaMalloc(&h->rows, sizeof(PointRow) * cs));
  CK(cudaHostAlloc(&h->h_parts, sizeof(RRPart) * ML_NPART_MAX, cudaHostAllocMapped));
  memset(h->h_parts, 0, sizeof(RRPart) * ML_NPART_MAX);
  *out = h;
  return ML_OK;
}

void ml_destroy(ml_handle *h) {
  if (!h)
    return;
  use_device(h);
  cudaStreamSynchronize(h->stream);
  if (h->host_prof && h->host_prof_n)
    fprintf(stderr, "[mins host prof] %ld one-kernel scan updates: consts %.1f us, resolve+launch %.1f us, wait+host tail %.1f us\n",
            h->host_prof_n, h->host_prof_us[0] / h->host_prof_n, h->host_prof_us[1] / h->host_prof_n, h->host_prof_us[2] / h->host_prof_n);
  if (h->h_merge)
    cudaFreeHost(h->h_merge);
  void *dev[] = {h->map_pts[0], h->map_pts[1], h->d_map_n, h->d_grid, h->d_bbox, h->cell_cnt, h->cell_start, h->cell_pts, h->pt_cell,
                 h->pt_rank, h->tile_sums, h->vox_tab, h->add_slot, h->contender, h->keep, h->pos, h->d_ints, h->add_pts, h->d_raw,
                 h->scan_pts, h->scan_orig, h->scan_map, h->scan_tmp, h->d_st, h->sort_keys[0], h->sort_keys[1], h->sort_vals[0],
                 h->sort_vals[1], h->radix_state, h->radix_ghist, h->d_vg, h->d_knots, h->qtab, h->q_slot, h->q_rank, h->q_sorted, h->units,
                 h->knn_idx, h->knn_d2, h->knn_ok, h->nbr, h->d_T, h->d_fctl, h->units_tag, h->rows,
                 h->sim_planes, h->sim_dirs, h->sim_dist, h->sim_n_out, h->icp_keep, h->icp_normals, h->icp_cell_pts, h->icp_read,
                 h->icp_q, h->d_icp, h->d_ist, h->icp_parts, h->d_hrows, h->tl_rows, h->lb_state, h->lb_ctr, h->radix_dbase, h->sim_knots, h->mg_done, h->d_bbox_vg};
  for (void *p : dev)
    if (p)
      cudaFree(p);
  if (h->h_raw)
    cudaFreeHost(h->h_raw);
  if (h->h_parts)
    cudaFreeHost(h->h_parts);
  if (h->h_icp)
    cudaFreeHost(h->h_icp);
  for (auto &e : h->ev)
    cudaEventDestroy(e);
  cudaStreamDestroy(h->stream);
  delete h;
}

// ---------------------------------------------------------------------------------------------- map
int ml_map_reset(ml_handle *h) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  h->rows_valid = false;
  h->map_n = 0;
  h->map_init = false;
  CK(cudaMemsetAsync(h->d_map_n, 0, sizeof(int), h->stream));
  CK(cudaMemsetAsync(h->d_grid, 0, sizeof(GridDesc), h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return ML_OK;
}
int ml_map_initialized(const ml_handle *h) {
  if (!h)
    return 0;
  resolve_merge(const_cast<ml_handle *>(h)); // a failed merge leaves the previous map in place: still the right answer
  return h->map_init ? 1 : 0;
}

int ml_map_build(ml_handle *h, const float *xyzi, int n) {
  if (!h || n < 0 || (n > 0 && !xyzi))
    return fail(h, ML_ERR_ARG, "ml_map_build: bad argument");
  RESOLVE(h);
  h->rows_valid = false;
  if (n > h->cap_map)
    return fail(h, ML_ERR_CAPACITY, "ml_map_build: more points than max_map_points");
  h->cur = 0;
  if (n > 0)
    CK(cudaMemcpyAsync(h->map_pts[0], xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, h->stream));
  LAUNCH(k_set_int, 1, 1, h->d_map_n, n);
  h->map_n = n;
  return rebuild_grid(h);
}

int ml_map_build_from_scan(ml_handle *h) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  h->rows_valid = false;
  ScanState st;
  CK(cudaMemcpyAsync(&st, h->d_st, sizeof(st), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (st.n_scan > h->cap_map)
    return fail(h, ML_ERR_CAPACITY, "ml_map_build_from_scan: scan larger than max_map_points");
  h->cur = 0;
  if (st.n_scan > 0)
    CK(cudaMemcpyAsync(h->map_pts[0], h->scan_pts, sizeof(float4) * (size_t)st.n_scan, cudaMemcpyDeviceToDevice, h->stream));
  LAUNCH(k_set_int, 1, 1, h->d_map_n, st.n_scan);
  h->map_n = st.n_scan;
  return rebuild_grid(h);
}

// shared tail of ml_map_add / ml_map_add_scan: *d_nadd new points (host upper bound n_ub) into the map. xf_src != NULL: the new
// points are xf_src[] in the sensor frame, transformed by T_LtoM into add_pts on the way (ml_map_add_scan); else add_pts holds them.
static int map_add_device(ml_handle *h, const float4 *xf_src, const int *d_nadd, int n_ub, int downsample, float ds, int *n_live) {
  h->rows_valid = false; // the map changes: a later ml_point_rows must not re-run the rows kernel against it
  if (n_ub <= 0) {
    if (n_live)
      *n_live = h->map_n;
    return ML_OK;
  }
  if (downsample && !(ds > 0.f))
    return fail(h, ML_ERR_ARG, "ml_map_add: downsample size must be positive");
  const int M = h->map_n;
  int rc;
  if (downsample) {
    const uint32_t mask = (uint32_t)h->vox_slots - 1u;
    LAUNCH(k_vox_clear, std::min(148 * 8, cdiv(h->vox_slots, 256)), 256, h->vox_tab, h->vox_slots);
    if (xf_src)
      LAUNCH(k_vox_insert_new<true>, cdiv(n_ub, 256), 256, xf_src, h->T_LtoM, h->add_pts, d_nadd, h->d_map_n, ds, h->vox_tab, mask, h->add_slot);
    else
      LAUNCH(k_vox_insert_new<false>, cdiv(n_ub, 256), 256, xf_src, h->T_LtoM, h->add_pts, d_nadd, h->d_map_n, ds, h->vox_tab, mask, h->add_slot);
    if (M > 0)
      LAUNCH(k_vox_compete_old, cdiv(M, 256), 256, h->map_pts[h->cur], h->d_map_n, ds, h->vox_tab, mask, h->contender);
    rc = merge_into_map<MG_VOX>(h, h->map_pts[h->cur], h->d_map_n, h->add_pts, d_nadd, M + n_ub, "ml_map_add", n_live != nullptr);
  } else {
    if (xf_src)
      LAUNCH(k_transform, cdiv(n_ub, 256), 256, xf_src, d_nadd, h->T_LtoM, h->add_pts);
    rc = merge_into_map<MG_ALL>(h, h->map_pts[h->cur], h->d_map_n, h->add_pts, d_nadd, M + n_ub, "ml_map_add", n_live != nullptr);
  }
  if (n_live)
    *n_live = h->map_n;
  return rc;
}

int ml_map_add(ml_handle *h, const float *xyzi, int n, int downsample, float ds, int *n_live) {
  if (!h || n < 0 || (n > 0 && !xyzi))
    return fail(h, ML_ERR_ARG, "ml_map_add: bad argument");
  RESOLVE(h);
  if (n > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_map_add: more points than max_scan_points in one call");
  if (n > 0)
    CK(cudaMemcpyAsync(h->add_pts, xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, h->stream));
  LAUNCH(k_set_int, 1, 1, h->d_ints + 0, n);
  return map_add_device(h, nullptr, h->d_ints + 0, n, downsample, ds, n_live);
}

int ml_map_add_scan(ml_handle *h, int downsample, float ds, int *n_live) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  if (!h->have_T)
    return fail(h, ML_ERR_ARG, "ml_map_add_scan: ml_scan_transform has not been called for this scan");
  return map_add_device(h, h->scan_orig, ST(n_orig), h->n_raw, downsample, ds, n_live);
}

// voxel-grid centroid filter of src[0..*d_n) -> dst, *d_n updated (scan.cuh: 7 launches). n_ub host bound. tmp buffers: sort_*.
// d_vox_n / d_vox_pass (may be NULL): occupied voxels and PCL's "leaf too small" verdict, for the scan state.
// bin: the output is the scan's query cloud - also transform it by h->T_LtoM into scan_map and bin it for the kNN.
static int voxel_filter(ml_handle *h, const float4 *src, float4 *dst, int *d_n, int n_ub, float leaf, int *d_vox_n, int *d_vox_pass, bool bin,
                        bool have_vg = false) {
  if (n_ub <= 0)
    return ML_OK;
  if (!have_vg)
    LAUNCH(k_bbox_voxgrid, std::min(148 * 8, cdiv(n_ub, 256)), 256, src, d_n, h->d_bbox_vg, h->mg_done, leaf, h->d_vg);
  LAUNCH(k_vox_index_hist, cdiv(n_ub, RS_TILE), RS_BLOCK, src, d_n, h->d_vg, h->sort_keys[0], h->sort_vals[0], h->radix_ghist,
         h->radix_dbase, h->mg_done);
  radix_sort(h, d_n, n_ub, 32);
  const uint32_t mask = (uint32_t)h->q_slots - 1u;
  if (bin)
    LAUNCH(k_vox_reduce<true>, cdiv(n_ub, VR_TILE), VR_BLOCK, src, h->sort_keys[0], h->sort_vals[0], d_n, h->d_vg, dst, h->lb_state,
           h->lb_ctr, lb_next_epoch(h), h->mg_done, h->d_ints + 6, d_vox_n, d_vox_pass, h->T_LtoM, h->scan_map, h->d_grid, h->qtab, mask,
           h->q_slot, h->q_rank);
  else
    LAUNCH(k_vox_reduce<false>, cdiv(n_ub, VR_TILE), VR_BLOCK, src, h->sort_keys[0], h->sort_vals[0], d_n, h->d_vg, dst, h->lb_state,
           h->lb_ctr, lb_next_epoch(h), h->mg_done, h->d_ints + 6, d_vox_n, d_vox_pass, h->T_LtoM, (float4 *)nullptr,
           (const GridDesc *)nullptr, (QSlot *)nullptr, mask, (uint32_t *)nullptr, (uint32_t *)nullptr);
  return check_launch(h);
}

int ml_map_reanchor(ml_handle *h, const double T[16], double now_minus_FT, double decay_time, double decay_dist, int do_downsample,
                    float ds, int *n_live) {
  if (!h || !T)
    return fail(h, ML_ERR_ARG, "ml_map_reanchor: bad argument");
  RESOLVE(h);
  h->rows_valid = false;
  if (!h->map_init)
    return fail(h, ML_SOFT_NO_MAP, "ml_map_reanchor: map not initialised");
  const int M = h->map_n;
  const float4 *src = h->map_pts[h->cur];
  int *d_n = h->d_ints + 5;
  LAUNCH(k_copy_int, 1, 1, d_n, h->d_map_n);
  if (do_downsample) {
    int rc = voxel_filter(h, src, h->scan_tmp, d_n, M, ds, nullptr, nullptr, false);
    if (rc)
      return rc;
    src = h->scan_tmp;
  }
  ReanchorArgs a;
  memcpy(a.T, T, sizeof(a.T));
  a.now_minus_FT = now_minus_FT, a.decay_time = decay_time, a.decay_dist = decay_dist;
  // transformed copy into cell_pts (free scratch: the grid is rebuilt below) + keep flags, then the one-pass merge
  LAUNCH(k_reanchor, cdiv(M, 256), 256, src, d_n, a, h->cell_pts, h->keep);
  int rc = merge_into_map<MG_FLAGS>(h, h->cell_pts, d_n, (const float4 *)nullptr, (const int *)nullptr, M, "ml_map_reanchor", n_live != nullptr);
  if (n_live)
    *n_live = h->map_n;
  return rc;
}

int ml_map_flatten(ml_handle *h, float *xyzi_out, int cap, int *n) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  int m = std::min(cap, h->map_n);
  if (m > 0 && xyzi_out) {
    CK(cudaMemcpyAsync(xyzi_out, h->map_pts[h->cur], sizeof(float4) * (size_t)m, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  if (n)
    *n = h->map_n;
  return ML_OK;
}
int ml_map_size(ml_handle *h, int *n) {
  if (!h || !n)
    return ML_ERR_ARG;
  RESOLVE(h);
  *n = h->map_n;
  return ML_OK;
}

// ---------------------------------------------------------------------------------------------- scan
// Where the kernels read the raw scan from. *over_pcie: the kernel itself pulls the bytes across the bus.
//  * device (or managed) memory: used in place, no copy (e.g. the simulator's ray-caster, a device-side driver)
//  * pinned, device-mapped host memory and the one-pass front end: the kernel reads the scan straight over PCIe (each point
//    is read exactly once), so there is no separate DMA copy to wait for
//  * pinned without a device mapping: DMA copy; pageable: the kernel reads the handle's own pinned staging buffer, which the
//    host fills with one memcpy (no DMA node either: the copy engine's launch latency is longer than the memcpy)
static int resolve_input(ml_handle *h, const float *xyz, int stride, int n, const float **raw_src, bool *over_pcie) {
  const size_t bytes = sizeof(float) * (size_t)stride * (size_t)n;
  cudaPointerAttributes at;
  const bool known = cudaPointerGetAttributes(&at, xyz) == cudaSuccess;
  cudaGetLastError();
  const bool pinned = known && at.type == cudaMemoryTypeHost;
  const bool on_device = known && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
  *over_pcie = false;
  if (on_device) {
    *raw_src = xyz;
  } else if (pinned && h->zero_copy && at.devicePointer && h->use_fused_front) {
    *raw_src = (const float *)at.devicePointer;
    *over_pcie = true;
  } else if (!pinned && h->zero_copy && h->use_fused_front && h->h_raw_dev) {
    // (widening packed triples to PointXYZ here, so that the kernel could take its 16-byte path, was measured: the scalar repack
    // costs more than the kernel gains, 288 us against 255 us per 131 072-point scan)
    memcpy(h->h_raw, xyz, bytes);
    *raw_src = h->h_raw_dev;
    *over_pcie = true;
  } else {
    const float *src = xyz;
    if (!pinned) { // pageable host memory: stage through the handle's pinned buffer
      memcpy(h->h_raw, xyz, bytes);
      src = h->h_raw;
    }
    CK(cudaMemcpyAsync(h->d_raw, src, bytes, cudaMemcpyHostToDevice, h->stream));
    *raw_src = h->d_raw;
  }
  return ML_OK;
}

// LiDARData ctor on the device. T_LtoM (may be NULL): when the map transform is already known the compaction pass writes
// pointcloud_in_map as well (LidarHelper.cpp:192), saving one pass over the scan.
// vg_leaf > 0: a voxel filter with this leaf follows at once (one-call path): the front end also prepares its grid (h->vg_ready)
static int scan_upload_async(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, const float *T_LtoM,
                             float vg_leaf = 0.f) {
  if (n < 0 || (n > 0 && !xyz) || (stride != 3 && stride != 4))
    return fail(h, ML_ERR_ARG, "ml_scan_upload: bad argument (stride must be 3 or 4)");
  if (n > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_scan_upload: more points than max_scan_points");
  drop_stale_bins(h);
  h->n_raw = n;
  h->have_T = false;
  h->rows_valid = false;
  h->vg_ready = false;
  if (T_LtoM) {
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 4; c++)
        h->T_LtoM.m[4 * r + c] = T_LtoM[4 * r + c];
    h->have_T = true;
  }
  if (n == 0) {
    LAUNCH(k_scanstate_zero, 1, 1, h->d_st);
    return check_launch(h);
  }
  const float *raw_src = nullptr;
  bool over_pcie = false;
  if (int rc_in = resolve_input(h, xyz, stride, n, &raw_src, &over_pcie))
    return rc_in;
  const int tiles = cdiv(n, RF_TILE);
  float4 *in_map = T_LtoM ? h->scan_map : nullptr;
  if (h->use_fused_front) {
    // one launch: range filter (single-pass stable compaction) [+ transform + query binning when the map transform is known]
    const uint32_t epoch = lb_next_epoch(h);
    const uint32_t mask = (uint32_t)h->q_slots - 1u;
    const bool vec4 = stride == 4 && (((uintptr_t)raw_src) & 15u) == 0;
    const bool bin = T_LtoM && h->map_init;
#define FRONT(V, B)                                                                                                    \
  LAUNCHP((k_range_filter_bin<V, B>), tiles, RF_BLOCK, raw_src, stride, n, t_rel, (double)rmin, (double)rmax, h->d_st, h->scan_pts,  \
          h->scan_orig, h->T_LtoM, in_map, h->lb_state, h->lb_ctr, epoch, h->d_grid, h->qtab, mask, h->q_slot, h->q_rank, bin ? 0.f : vg_leaf,        \
          h->d_bbox_vg, h->mg_done, h->d_vg)
    if (vec4 && bin)
      FRONT(true, true);
    else if (vec4)
      FRONT(true, false);
    else if (bin)
      FRONT(false, true);
    else
      FRONT(false, false);
#undef FRONT
    h->bin_done = bin;
    h->vg_ready = !bin && vg_leaf > 0.f;
    return check_launch(h);
  }
  if (stride == 4 && (((uintptr_t)raw_src) & 15u) == 0) {
    LAUNCHP(k_range_count<true>, tiles, RF_BLOCK, raw_src, stride, n, (double)rmin, (double)rmax, h->d_st, h->tile_sums);
    LAUNCHP(k_range_compact<true>, tiles, RF_BLOCK, raw_src, stride, n, t_rel, (double)rmin, (double)rmax, h->tile_sums, h->d_st,
           h->scan_pts, h->scan_orig, h->T_LtoM, in_map);
  } else {
    LAUNCHP(k_range_count<false>, tiles, RF_BLOCK, raw_src, stride, n, (double)rmin, (double)rmax, h->d_st, h->tile_sums);
    LAUNCHP(k_range_compact<false>, tiles, RF_BLOCK, raw_src, stride, n, t_rel, (double)rmin, (double)rmax, h->tile_sums, h->d_st,
           h->scan_pts, h->scan_orig, h->T_LtoM, in_map);
  }
  return check_launch(h);
}

int ml_scan_upload(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, int *n_kept) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  int rc = scan_upload_async(h, xyz, stride, n, t_rel, rmin, rmax, nullptr);
  if (rc)
    return rc;
  if (n_kept) {
    ScanState st;
    CK(cudaMemcpyAsync(&st, h->d_st, sizeof(st), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *n_kept = st.n_orig;
  }
  return ML_OK;
}

void *ml_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (bytes == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void ml_host_free(void *p) {
  if (p)
    cudaFreeHost(p);
}

int ml_scan_set(ml_handle *h, const float *xyzi, int n) {
  if (!h || n < 0 || (n > 0 && !xyzi))
    return fail(h, ML_ERR_ARG, "ml_scan_set: bad argument");
  use_device(h);
  if (n > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_scan_set: more points than max_scan_points");
  drop_stale_bins(h);
  h->n_raw = n;
  h->have_T = false;
  h->rows_valid = false;
  h->vg_ready = false;
  ScanState st;
  memset(&st, 0, sizeof(st));
  st.n_raw = st.n_orig = st.n_scan = n;
  if (n > 0) {
    CK(cudaMemcpyAsync(h->scan_pts, xyzi, sizeof(float4) * (size_t)n, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->scan_orig, h->scan_pts, sizeof(float4) * (size_t)n, cudaMemcpyDeviceToDevice, h->stream));
  }
  CK(cudaMemcpyAsync(h->d_st, &st, sizeof(st), cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  return ML_OK;
}

int ml_scan_deskew(ml_handle *h, const float *T_LitoL_knots, int n_knots, int total_points) {
  if (!h || !T_LitoL_knots || n_knots < 2 || total_points < 1)
    return fail(h, ML_ERR_ARG, "ml_scan_deskew: bad argument");
  use_device(h);
  h->rows_valid = false;
  const size_t need = (size_t)n_knots * 12;
  if (need > h->knots_cap) {
    if (h->d_knots)
      cudaFree(h->d_knots);
    CK(cudaMalloc(&h->d_knots, sizeof(float) * need));
    h->knots_cap = need;
  }
  CK(cudaMemcpyAsync(h->d_knots, T_LitoL_knots, sizeof(float) * need, cudaMemcpyHostToDevice, h->stream));
  if (h->n_raw > 0)
    LAUNCH(k_deskew, cdiv(h->n_raw, 256), 256, h->scan_pts, ST(n_scan), h->d_knots, n_knots, total_points);
  CK(cudaStreamSynchronize(h->stream));
  return check_launch(h);
}

// T_LtoM != NULL (one-call path): the map transform is already known, so the filter's last kernel also produces
// pointcloud_in_map and bins the queries (what ml_scan_transform and the binning launch would do next)
static int scan_downsample_async(ml_handle *h, float leaf, const float *T_LtoM) {
  if (!(leaf > 0.f))
    return fail(h, ML_ERR_ARG, "ml_scan_downsample: leaf must be positive");
  const bool bin = T_LtoM && h->map_init && h->use_fused_front && h->n_raw > 0;
  drop_stale_bins(h);
  if (bin) {
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 4; c++)
        h->T_LtoM.m[4 * r + c] = T_LtoM[4 * r + c];
  }
  int rc = voxel_filter(h, h->scan_pts, h->scan_tmp, ST(n_scan), h->n_raw, leaf, ST(vox_n), ST(vox_passthrough), bin, h->vg_ready);
  h->vg_ready = false;
  if (rc)
    return rc;
  std::swap(h->scan_pts, h->scan_tmp);
  h->rows_valid = false;
  if (bin) {
    h->have_T = true;
    h->bin_done = true;
  }
  return ML_OK;
}

int ml_scan_downsample(ml_handle *h, float leaf, int *n_ds) {
  if (!h)
    return ML_ERR_ARG;
  use_device(h);
  int rc = scan_downsample_async(h, leaf, nullptr);
  if (rc)
    return rc;
  if (n_ds) {
    ScanState st;
    CK(cudaMemcpyAsync(&st, h->d_st, sizeof(st), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    *n_ds = st.n_scan;
  }
  return ML_OK;
}

static int scan_transform_async(ml_handle *h, const float T[16]) {
  drop_stale_bins(h);
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++)
      h->T_LtoM.m[4 * r + c] = T[4 * r + c];
  h->have_T = true;
  h->rows_valid = false;
  if (h->n_raw > 0)
    LAUNCH(k_transform, cdiv(h->n_raw, 256), 256, h->scan_pts, ST(n_scan), h->T_LtoM, h->scan_map);
  return check_launch(h);
}

int ml_scan_transform(ml_handle *h, const float T_LtoM[16]) {
  if (!h || !T_LtoM)
    return fail(h, ML_ERR_ARG, "ml_scan_transform: bad argument");
  use_device(h);
  return scan_transform_async(h, T_LtoM);
}

int ml_scan_download(ml_handle *h, int which, float *xyzi_out, int cap, int *n) {
  if (!h || which < 0 || which > 2)
    return fail(h, ML_ERR_ARG, "ml_scan_download: bad argument");
  use_device(h);
  ScanState st;
  CK(cudaMemcpyAsync(&st, h->d_st, sizeof(st), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  int cnt = which == 2 ? st.n_orig : st.n_scan;
  const float4 *src = which == 0 ? h->scan_pts : (which == 1 ? h->scan_map : h->scan_orig);
  int m = std::min(cap, cnt);
  if (m > 0 && xyzi_out) {
    CK(cudaMemcpyAsync(xyzi_out, src, sizeof(float4) * (size_t)m, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  if (n)
    *n = cnt;
  return ML_OK;
}

// ---------------------------------------------------------------------------------------- simulator
int ml_sim_set_world(ml_handle *h, const double *planes, int n_planes) {
  if (!h || !planes || n_planes < 1)
    return fail(h, ML_ERR_ARG, "ml_sim_set_world: bad argument");
  use_device(h);
  CK(cudaStreamSynchronize(h->stream));
  if (h->sim_planes)
    cudaFree(h->sim_planes);
  h->sim_planes = nullptr;
  const size_t bytes = sizeof(double) * SIM_PLANE_STRIDE * (size_t)n_planes;
  CK(cudaMalloc(&h->sim_planes, bytes));
  CK(cudaMemcpyAsync(h->sim_planes, planes, bytes, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->sim_n_planes = n_planes;
  return ML_OK;
}

int ml_sim_set_rays(ml_handle *h, const double *dirs_L, int n_rays) {
  if (!h || !dirs_L || n_rays < 1)
    return fail(h, ML_ERR_ARG, "ml_sim_set_rays: bad argument");
  use_device(h);
  if (n_rays > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_sim_set_rays: more rays than max_scan_points");
  CK(cudaStreamSynchronize(h->stream));
  if (h->sim_dirs)
    cudaFree(h->sim_dirs);
  if (h->sim_dist)
    cudaFree(h->sim_dist);
  h->sim_dirs = h->sim_dist = nullptr;
  CK(cudaMalloc(&h->sim_dirs, sizeof(double) * 3 * (size_t)n_rays));
  CK(cudaMalloc(&h->sim_dist, sizeof(double) * (size_t)n_rays));
  if (!h->sim_n_out)
    CK(cudaMalloc(&h->sim_n_out, sizeof(int)));
  CK(cudaMemcpyAsync(h->sim_dirs, dirs_L, sizeof(double) * 3 * (size_t)n_rays, cudaMemcpyHostToDevice, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  h->sim_n_rays = n_rays;
  return ML_OK;
}

static int sim_raycast_common(ml_handle *h, const SimPose &pose, const double *d_knots, int n_knots, double min_range, double max_range,
                              float noise, uint64_t seed, float *xyz_out, int cap, int *n_out) {
  const int n = h->sim_n_rays, tiles = cdiv(n, RF_TILE);
  if (d_knots)
    LAUNCH(k_ray_cast<true>, tiles, SIM_BLOCK, h->sim_planes, h->sim_n_planes, h->sim_dirs, n, pose, d_knots, n_knots, min_range, max_range,
           h->sim_dist, h->tile_sums);
  else
    LAUNCH(k_ray_cast<false>, tiles, SIM_BLOCK, h->sim_planes, h->sim_n_planes, h->sim_dirs, n, pose, d_knots, n_knots, min_range, max_range,
           h->sim_dist, h->tile_sums);
  LAUNCH(k_ray_emit, tiles, SIM_BLOCK, h->sim_dist, h->sim_dirs, n, h->tile_sums, noise, (uint32_t)(seed & 0xffffffffu),
         (uint32_t)(seed >> 32), reinterpret_cast<float4 *>(h->d_raw), h->sim_n_out);
  CK(cudaMemcpyAsync(&h->sim_last_n, h->sim_n_out, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaGetLastError());
  const int m = std::min(cap, h->sim_last_n);
  if (xyz_out && m > 0) {
    CK(cudaMemcpyAsync(xyz_out, h->d_raw, sizeof(float) * 4 * (size_t)m, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  if (n_out)
    *n_out = h->sim_last_n;
  return ML_OK;
}

int ml_sim_raycast(ml_handle *h, const double p_LinG[3], const double R_LtoG[9], double min_range, double max_range, float noise,
                   uint64_t seed, float *xyz_out, int cap, int *n_out) {
  if (!h || !p_LinG || !R_LtoG)
    return fail(h, ML_ERR_ARG, "ml_sim_raycast: bad argument");
  use_device(h);
  if (!h->sim_planes || !h->sim_dirs)
    return fail(h, ML_ERR_ARG, "ml_sim_raycast: ml_sim_set_world / ml_sim_set_rays have not been called");
  SimPose pose;
  for (int a = 0; a < 3; a++)
    pose.o[a] = p_LinG[a];
  for (int a = 0; a < 9; a++)
    pose.R[a] = R_LtoG[a];
  return sim_raycast_common(h, pose, nullptr, 0, min_range, max_range, noise, seed, xyz_out, cap, n_out);
}

int ml_sim_raycast_sweep(ml_handle *h, const double *pose_knots, int n_knots, double min_range, double max_range, float noise,
                         uint64_t seed, float *xyz_out, int cap, int *n_out) {
  if (!h || !pose_knots || n_knots < 2)
    return fail(h, ML_ERR_ARG, "ml_sim_raycast_sweep: bad argument (at least 2 pose knots)");
  use_device(h);
  if (!h->sim_planes || !h->sim_dirs)
    return fail(h, ML_ERR_ARG, "ml_sim_raycast_sweep: ml_sim_set_world / ml_sim_set_rays have not been called");
  const size_t need = (size_t)n_knots * 12;
  if (need > h->sim_knots_cap) {
    CK(cudaStreamSynchronize(h->stream));
    if (h->sim_knots)
      cudaFree(h->sim_knots);
    h->sim_knots = nullptr;
    CK(cudaMalloc(&h->sim_knots, sizeof(double) * need));
    h->sim_knots_cap = need;
  }
  CK(cudaMemcpyAsync(h->sim_knots, pose_knots, sizeof(double) * need, cudaMemcpyHostToDevice, h->stream));
  SimPose pose;
  memset(&pose, 0, sizeof(pose));
  return sim_raycast_common(h, pose, h->sim_knots, n_knots, min_range, max_range, noise, seed, xyz_out, cap, n_out);
}

int ml_sim_scan_device(ml_handle *h, const float **xyz_dev, int *n) {
  if (!h || !xyz_dev || !n)
    return fail(h, ML_ERR_ARG, "ml_sim_scan_device: bad argument");
  *xyz_dev = h->d_raw;
  *n = h->sim_last_n;
  return ML_OK;
}

// -------------------------------------------------------------------------------------------- update
} // extern "C"

template <int K> static void launch_knn(ml_handle *h, float max_d2, int ph) {
  const int blocks = h->n_sm * 8;
  LAUNCHP(k_knn<K>, blocks, KNN_WARPS * 32, h->units, h->d_st, ph, h->q_sorted, h->d_grid, h->cell_start, h->cell_pts, max_d2,
         h->knn_idx, h->knn_d2, h->knn_ok, h->qtab, h->q_slots, (const float4 *)h->map_pts[h->cur], h->nbr,
         h->tl_on ? h->tl_knn : (unsigned long long *)nullptr);
}
// fused rows + reduction + finalize; rows_out != nullptr only for the parity export
template <int K, int NB, int BLOCK, bool EXPORT> static int launch_rows_cfg(ml_handle *h, PointRow *rows_out, int nb) {
  constexpr int WARPS = BLOCK / 32;
  const size_t smem = RRLayout<NB>::smem(WARPS);
  // the attribute is per device (and per function): remembered per handle, a handle lives on one device
  bool &attr_set = h->rows_attr[K][(NB == 5 ? 1 : 0) + (BLOCK == RR_BLOCK_BIG ? 2 : BLOCK == RR_BLOCK_MID ? 4 : 0) + (EXPORT ? 6 : 0)];
  if (!attr_set) {
    CK(cudaFuncSetAttribute(k_rows_reduce<K, NB, BLOCK, EXPORT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  h->rows_nb = nb;
  h->parts_n = nb;
  h->parts_armed = 0; // this kernel overwrites the parts: nothing is armed afterwards
  launch_pdl(h, k_rows_reduce<K, NB, BLOCK, EXPORT>, nb, BLOCK, smem, h->rowc, (const ScanState *)h->d_st, h->n_raw,
             (const float4 *)h->scan_pts, (const float4 *)h->nbr, rows_out, h->h_parts /* pinned host memory */,
             h->tl_on ? h->tl_rows : (unsigned long long *)nullptr);
  return ML_OK;
}
// Launch shape: one CTA per SM (fewer for small scans), and the block size whose warps walk the 32-point tiles in the most even
// number of rounds - the kernel ends with a CTA barrier, so a round that only some warps run is paid by all of them. Fixed for a
// given scan size: the reduction order (and the result, bit for bit) depends on nothing else.
template <int K, int NB> static int launch_rows_nb(ml_handle *h, PointRow *rows_out) {
  const int n = std::max(h->n_raw, 1);
  const int ntiles = cdiv(n, 32);
  const int nb = std::max(1, std::min(h->n_sm, cdiv(n, 256)));
  if (rows_out) // parity export: same arithmetic, same launch-shape rule is not needed (the sums are not compared bit for bit
    return launch_rows_cfg<K, NB, RR_BLOCK, true>(h, rows_out, nb); // with the timed run's; the export re-writes the result block)
  if (NB == 3 && K == 5) {
    // efficiency of a block size = tiles / (rounds * warps); more warps also hide more latency, so they win ties
    auto eff = [&](int warps) { return (double)ntiles / ((double)cdiv(ntiles, nb * warps) * nb * warps); };
    int best = RR_BLOCK;
    double be = eff(RR_BLOCK / 32) * 0.92;
    for (int b : {RR_BLOCK_MID, RR_BLOCK_BIG})
      if (eff(b / 32) >= be)
        best = b, be = eff(b / 32);
    if (h->rows_block)
      best = h->rows_block;
    constexpr int MID = (NB == 3 && K == 5) ? RR_BLOCK_MID : RR_BLOCK, BIG = (NB == 3 && K == 5) ? RR_BLOCK_BIG : RR_BLOCK;
    if (best == RR_BLOCK_BIG)
      return launch_rows_cfg<K, NB, BIG, false>(h, rows_out, nb);
    if (best == RR_BLOCK_MID)
      return launch_rows_cfg<K, NB, MID, false>(h, rows_out, nb);
  }
  return launch_rows_cfg<K, NB, RR_BLOCK, false>(h, rows_out, nb);
}
// NB = 3-wide blocks of the folded Jacobian direction that can be non-zero: 3 without extrinsic calibration, 5 with
template <int K> static int launch_rows(ml_handle *h, PointRow *rows_out) {
  return h->consts.calib_col >= 0 ? launch_rows_nb<K, 5>(h, rows_out) : launch_rows_nb<K, 3>(h, rows_out);
}
static int launch_rows_k(ml_handle *h, int k, PointRow *rows_out) {
  switch (k) {
  case 3: return launch_rows<3>(h, rows_out);
  case 4: return launch_rows<4>(h, rows_out);
  case 5: return launch_rows<5>(h, rows_out);
  case 6: return launch_rows<6>(h, rows_out);
  case 7: return launch_rows<7>(h, rows_out);
  case 8: return launch_rows<8>(h, rows_out);
  case 9: return launch_rows<9>(h, rows_out);
  default: return launch_rows<10>(h, rows_out);
  }
}

static int knn_async(ml_handle *h, int k, float max_d2) {
  if (k < 1 || k > ML_MAX_K)
    return fail(h, ML_ERR_ARG, "knn: k out of range [1, ML_MAX_K]");
  if (!h->map_init)
    return fail(h, ML_SOFT_NO_MAP, "knn: map not initialised");
  if (!h->have_T)
    return fail(h, ML_ERR_ARG, "knn: ml_scan_transform has not been called for this scan");
  const int n = h->n_raw;
  if (n <= 0)
    return ML_OK;
  const uint32_t mask = (uint32_t)h->q_slots - 1u;
  const int ph = h->knn_calls & 1;
  h->knn_calls++;
  if (h->timing == 1)
    cudaEventRecord(h->ev[3], h->stream);
  if (!h->bin_done)
    LAUNCHP(k_query_bin, cdiv(n, 256), 256, h->scan_map, h->d_st, h->d_grid, h->qtab, mask, h->q_slot, h->q_rank);
  h->bin_done = false;
  LAUNCHP(k_unit_alloc, cdiv(h->q_slots, 256), 256, h->qtab, h->q_slots, h->d_st, ph, h->units, h->units_cap);
  LAUNCHP(k_query_scatter, cdiv(n, 256), 256, h->scan_map, h->d_st, h->qtab, h->q_slot, h->q_rank, h->q_sorted);
  if (h->timing == 1)
    cudaEventRecord(h->ev[4], h->stream);
  switch (k) {
  case 1: launch_knn<1>(h, max_d2, ph); break;
  case 2: launch_knn<2>(h, max_d2, ph); break;
  case 3: launch_knn<3>(h, max_d2, ph); break;
  case 4: launch_knn<4>(h, max_d2, ph); break;
  case 5: launch_knn<5>(h, max_d2, ph); break;
  case 6: launch_knn<6>(h, max_d2, ph); break;
  case 7: launch_knn<7>(h, max_d2, ph); break;
  case 8: launch_knn<8>(h, max_d2, ph); break;
  case 9: launch_knn<9>(h, max_d2, ph); break;
  default: launch_knn<10>(h, max_d2, ph); break;
  }
  if (h->timing == 1)
    cudaEventRecord(h->ev[5], h->stream);
  return check_launch(h);
}

extern "C" {

int ml_knn(ml_handle *h, int k, float max_d2, int32_t *idx, float *d2, uint8_t *ok, int *n) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  h->rows_valid = false; // the kNN buffers the rows kernel read are overwritten
  int rc = knn_async(h, k, max_d2);
  if (rc)
    return rc;
  ScanState st;
  CK(cudaMemcpyAsync(&st, h->d_st, sizeof(st), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  const size_t m = (size_t)st.n_scan;
  if (idx && m)
    CK(cudaMemcpyAsync(idx, h->knn_idx, sizeof(int32_t) * m * k, cudaMemcpyDeviceToHost, h->stream));
  if (d2 && m)
    CK(cudaMemcpyAsync(d2, h->knn_d2, sizeof(float) * m * k, cudaMemcpyDeviceToHost, h->stream));
  if (ok && m)
    CK(cudaMemcpyAsync(ok, h->knn_ok, m, cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (n)
    *n = st.n_scan;
  return ML_OK;
}

static int check_consts(ml_handle *h, const ml_scan_consts *c) {
  if (!c)
    return fail(h, ML_ERR_ARG, "linsys: null constants");
  if (c->total_hx < 1 || c->total_hx > ML_MAX_HX)
    return fail(h, ML_ERR_ARG, "linsys: total_hx out of range");
  if (c->map_ngbr_num < 3 || c->map_ngbr_num > ML_MAX_K)
    return fail(h, ML_ERR_ARG, "linsys: map_ngbr_num must be in [3, ML_MAX_K]");
  if (c->nI < 0 || c->nI > ML_MAX_BLOCKS || c->nA < 0 || c->nA > ML_MAX_BLOCKS)
    return fail(h, ML_ERR_ARG, "linsys: too many Jacobian blocks");
  for (int b = 0; b < c->nI + c->nA; b++) {
    const ml_jac_block &B = b < c->nI ? c->dTdxI[b] : c->dTdxA[b - c->nI];
    if (B.width < 1 || B.width > 6 || B.col < 0 || B.col + B.width > c->total_hx)
      return fail(h, ML_ERR_ARG, "linsys: Jacobian block outside [0, total_hx)");
  }
  if (c->calib_col >= 0 && c->calib_col + 6 > c->total_hx)
    return fail(h, ML_ERR_ARG, "linsys: calibration block outside [0, total_hx)");
  return ML_OK;
}

// 3x3 helpers (row-major) for the host-side constant folding
static inline void m3_mul(const double *A, const double *B, double *C) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
static inline void m3_vec(const double *A, const double *v, double *o) {
  for (int i = 0; i < 3; i++)
    o[i] = A[3 * i] * v[0] + A[3 * i + 1] * v[1] + A[3 * i + 2] * v[2];
}
// ml_scan_consts (the reference's local variables, UpdaterLidar.cpp:223-254) -> the folded form the rows kernel takes as its
// parameter (RowConsts, rows.cuh)
static void make_row_consts(const ml_scan_consts &c, RowConsts *r) {
  memcpy(r->RLtoM_est, c.RLtoM_est, sizeof r->RLtoM_est), memcpy(r->pLinM_est, c.pLinM_est, sizeof r->pLinM_est);
  memcpy(r->RLtoM_fej, c.RLtoM_fej, sizeof r->RLtoM_fej), memcpy(r->pLinM_fej, c.pLinM_fej, sizeof r->pLinM_fej);
  memcpy(r->RLtoI, c.RLtoI_fej, sizeof r->RLtoI);
  double t[3], u[3];
  m3_vec(c.RLtoI_fej, c.pIinL_fej, t); // w = RLtoI (pfinL - pIinL) = RLtoI pfinL - RLtoI pIinL
  for (int a = 0; a < 3; a++)
    r->pLinI[a] = -t[a];
  m3_mul(c.RGtoM_fej, c.RItoG_fej, r->RItoM);
  memcpy(r->RGtoM, c.RGtoM_fej, sizeof r->RGtoM);
  memcpy(r->RItoL, c.RItoL_fej, sizeof r->RItoL);
  m3_mul(c.RGtoA_fej, c.RLtoG_fej, r->U1);
  m3_vec(c.RItoG_fej, c.pLinI_fej, t);
  for (int a = 0; a < 3; a++)
    u[a] = c.pIinG_fej[a] + t[a] - c.pAinG_fej[a];
  m3_vec(c.RGtoA_fej, u, r->U0);
  for (int a = 0; a < 3; a++)
    u[a] = c.pIinG_fej[a] - c.pAinG_fej[a];
  m3_vec(c.RGtoM_fej, u, r->c1);
  memcpy(r->pIinL, c.pIinL_fej, sizeof r->pIinL);
  r->raw_noise = c.raw_noise, r->map_noise = c.map_noise;
  r->condi2 = c.plane_max_condi * c.plane_max_condi;
  r->p2pd = c.plane_max_p2pd;
  r->badmatch = c.plane_max_p2pd * 3.0;
  r->ie_o = c.ie_o, r->ie_p = c.ie_p;
  r->chi2_thr = c.chi2_mult * chi2_q95(c.map_ngbr_num - 2);
  r->intr = (!c.at_clone_m || !c.at_clone_a) ? 1 : 0;
  r->total_hx = c.total_hx;
}

// T (ML_G x ML_MAX_HX, row-major): h = T^T g. Rows = the 15 folded pose-space directions (ML_G, common.cuh), columns = state
// columns: the reference's dTdxI / dTdxA blocks at their map_hx columns (State.cpp:745-816, UpdaterLidar.cpp:343-347) and the
// identity on the calibration columns (:349). Columns of different blocks may coincide (I and A clones overlap): added in order.
static void build_T(const ml_scan_consts &c, double *T) {
  memset(T, 0, sizeof(double) * ML_G * ML_MAX_HX);
  for (int r = 0; r < 6; r++) // I rotation rows 0-2, I position rows 3-5
    for (int b = 0; b < c.nI; b++)
      for (int cc = 0; cc < c.dTdxI[b].width; cc++)
        T[r * ML_MAX_HX + c.dTdxI[b].col + cc] += c.dTdxI[b].D[6 * r + cc];
  for (int r = 3; r < 6; r++) // ... minus A position rows (g_A,pos = -g_I,pos)
    for (int b = 0; b < c.nA; b++)
      for (int cc = 0; cc < c.dTdxA[b].width; cc++)
        T[r * ML_MAX_HX + c.dTdxA[b].col + cc] -= c.dTdxA[b].D[6 * r + cc];
  for (int r = 6; r < 9; r++) // A rotation rows
    for (int b = 0; b < c.nA; b++)
      for (int cc = 0; cc < c.dTdxA[b].width; cc++)
        T[r * ML_MAX_HX + c.dTdxA[b].col + cc] += c.dTdxA[b].D[6 * (r - 6) + cc];
  if (c.calib_col >= 0)
    for (int r = 9; r < ML_G; r++)
      T[r * ML_MAX_HX + c.calib_col + (r - 9)] += 1.0;
}

// Per-scan constants, all on the host (~20 k multiply-adds, ~7 us): the folded rows parameter incl. Q = T P T^T (h^T P h = g^T Q g),
// and T for the host tail. On the two-kernel path they are computed while the device runs the front end and the kNN, so they cost
// the scan nothing; on the one-kernel path they are a parameter of the only kernel and precede its launch. There is no constants
// kernel, no side stream and no event anywhere.
static int consts_host(ml_handle *h, const ml_scan_consts *c, const double *P) {
  const int hx = c->total_hx;
  h->consts = *c;
  build_T(*c, h->hT);
  make_row_consts(*c, &h->rowc);
  const int NG = c->calib_col >= 0 ? ML_G : 9;
  double TP[ML_G][ML_MAX_HX], Q[ML_G][ML_G];
  int nz[ML_G][ML_MAX_HX], nnz[ML_G];
  for (int r = 0; r < NG; r++) {
    nnz[r] = 0;
    for (int l = 0; l < hx; l++)
      if (h->hT[r * ML_MAX_HX + l] != 0.0)
        nz[r][nnz[r]++] = l;
    for (int j = 0; j < hx; j++)
      TP[r][j] = 0.0;
    for (int q = 0; q < nnz[r]; q++) {
      const int l = nz[r][q];
      const double t = h->hT[r * ML_MAX_HX + l];
      const double *Pl = P + (size_t)l * hx;
      for (int j = 0; j < hx; j++)
        TP[r][j] += t * Pl[j];
    }
  }
  for (int r = 0; r < NG; r++)
    for (int q = 0; q < NG; q++) {
      double acc = 0.0;
      for (int e = 0; e < nnz[q]; e++)
        acc += TP[r][nz[q][e]] * h->hT[q * ML_MAX_HX + nz[q][e]];
      Q[r][q] = acc;
    }
  for (int a = 0; a < NG; a++)
    for (int b = 0; b < NG; b++)
      h->rowc.Qu[a * NG + b] = b < a ? 0.0 : (b == a ? Q[a][a] : Q[a][b] + Q[b][a]);
  return ML_OK;
}

// Host tail of the rows kernel: add the CTAs' reduced systems in CTA order (fixed: bit-reproducible), then the chain rule
// HtH = T^T G T, Htr = T^T b (T is block sparse: zero entries are skipped), and the counters.
typedef unsigned long long __attribute__((may_alias)) u64_alias; // the bit patterns of doubles
static int rows_host_tail(ml_handle *h, double *HtH, double *Htr, ml_counters *cnt, bool armed = false) {
  const ml_scan_consts &c = h->consts;
  const int hx = c.total_hx, nb = h->parts_n;
  const int NG = c.calib_col >= 0 ? ML_G : 9, NSYM = NG * (NG + 1) / 2, NE = NSYM + NG, NRED = NE + 4;
  double sum[ML_NRED];
  int cs[5] = {0, 0, 0, 0, 0};
  int nscan_max = -1;
  for (int e = 0; e < NRED; e++)
    sum[e] = 0.0;
  // armed (one-kernel path): the parts were armed before the launch; a word is taken as soon as it differs from the sentinel, in
  // CTA order all the same, so the sum runs beside the kernel's tail instead of after it, and the part is re-armed on the way.
  // A word that does not come within 20 ms means a faulted kernel: ask the stream.
  bool synced = false;
  const auto t0 = std::chrono::steady_clock::now();
  if (armed)
    h->parts_armed = 0; // until every part has been consumed and re-armed
  // a part is complete when none of its words is the sentinel any more (checked without an early exit: the loop vectorises)
  auto part_ready = [&](const RRPart &p) {
    const u64_alias *w = reinterpret_cast<const u64_alias *>(p.v);
    unsigned long long missing = 0;
    for (int e = 0; e < NRED; e++)
      missing += (w[e] == RR_ARMED_F64);
    for (int q = 0; q < 5; q++)
      missing += (p.c[q] == RR_ARMED_I32);
    missing += (p.n_scan == RR_ARMED_I32);
    return missing == 0;
  };
  for (int b = 0; b < nb; b++) {
    RRPart &p = h->h_parts[b];
    if (armed) {
      for (unsigned spins = 1; !part_ready(p); spins++) {
        asm volatile("" ::: "memory"); // re-read pinned memory
        if (synced)
          return fail(h, ML_ERR_CUDA, "ml_scan_update: the kernel finished without delivering its reduced system");
        if ((spins & 1023u) == 0u && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(20)) {
          CK(cudaStreamSynchronize(h->stream));
          CK(cudaGetLastError());
          synced = true;
        }
      }
      std::atomic_thread_fence(std::memory_order_acquire);
      for (int e = 0; e < NRED; e++)
        sum[e] += p.v[e];
      for (int q = 0; q < 5; q++)
        cs[q] += p.c[q];
      nscan_max = std::max(nscan_max, p.n_scan);
      u64_alias *w = reinterpret_cast<u64_alias *>(p.v);
      for (int e = 0; e < NRED; e++)
        w[e] = RR_ARMED_F64;
      for (int q = 0; q < 5; q++)
        p.c[q] = RR_ARMED_I32;
      p.n_scan = RR_ARMED_I32;
      continue;
    }
    for (int e = 0; e < NRED; e++)
      sum[e] += p.v[e];
    for (int q = 0; q < 5; q++)
      cs[q] += p.c[q];
  }
  if (armed) {
    h->last_n_scan = nscan_max;
    h->parts_armed = nb;
    h->parts_armed_nred = NRED;
  } else {
    h->last_n_scan = h->h_parts[0].n_scan;
  }
  if (cnt) {
    cnt->n_points = h->last_n_scan;
    cnt->n_acc = cs[0], cnt->n_knn_fail = cs[1], cnt->n_plane_fail = cs[2], cnt->n_badmatch = cs[3], cnt->n_chi_rej = cs[4];
    cnt->sum_chi = sum[NE + 0], cnt->sum_chi2 = sum[NE + 1], cnt->sum_res = sum[NE + 3], cnt->sum_res2 = sum[NE + 2];
  }
  if (HtH || Htr) {
    double G[ML_G][ML_G], GT[ML_G][ML_MAX_HX];
    {
      int e = 0;
      for (int a = 0; a < NG; a++)
        for (int b = a; b < NG; b++, e++)
          G[a][b] = G[b][a] = sum[e];
    }
    const double *T = h->hT;
    for (int r = 0; r < NG; r++) {
      for (int j = 0; j < hx; j++)
        GT[r][j] = 0.0;
      for (int l = 0; l < NG; l++) {
        const double gv = G[r][l];
        const double *Tl = T + l * ML_MAX_HX;
        for (int j = 0; j < hx; j++)
          GT[r][j] += gv * Tl[j];
      }
    }
    if (HtH) {
      for (int a = 0; a < hx; a++) {
        double *Ha = HtH + (size_t)a * hx;
        for (int b = a; b < hx; b++)
          Ha[b] = 0.0;
        for (int l = 0; l < NG; l++) {
          const double t = T[l * ML_MAX_HX + a];
          if (t == 0.0)
            continue;
          const double *Gl = GT[l];
          for (int b = a; b < hx; b++)
            Ha[b] += t * Gl[b];
        }
      }
      for (int a = 0; a < hx; a++) // upper triangle computed once and mirrored: exactly symmetric
        for (int b = 0; b < a; b++)
          HtH[(size_t)a * hx + b] = HtH[(size_t)b * hx + a];
    }
    if (Htr)
      for (int a = 0; a < hx; a++) {
        double acc = 0.0;
        for (int l = 0; l < NG; l++)
          acc += T[l * ML_MAX_HX + a] * sum[NSYM + l];
        Htr[a] = acc;
      }
  }
  return cs[0] > 0 ? ML_OK : ML_SOFT_NO_ROWS;
}

// rows kernel + the wait + the host tail; the kNN results of the resident scan must be in flight on the stream
static int rows_and_wait(ml_handle *h, double *HtH, double *Htr, ml_counters *cnt) {
  const int k = h->consts.map_ngbr_num;
  // residual statistics on (ml_set_stats): the variant that also runs the reference's Givens sequence per point and keeps the
  // per-point rows; otherwise the lean one
  int rc = launch_rows_k(h, k, h->stats_on ? h->rows : nullptr);
  if (rc)
    return rc;
  h->rows_exported = h->stats_on;
  if (h->timing == 1)
    cudaEventRecord(h->ev[6], h->stream);
  // no D2H copy: every CTA wrote its reduced system straight into the pinned result area
  if (h->timing)
    cudaEventRecord(h->ev[8], h->stream);
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaGetLastError());
  h->rows_valid = true;
  h->rows_need_knn = false;
  h->rows_k = k;
  return rows_host_tail(h, HtH, Htr, cnt);
}

// everything after the transform; one synchronisation at the end. consts_host must have been called.
static int linsys_async_and_wait(ml_handle *h, const ml_scan_consts *c, double *HtH, double *Htr, ml_counters *cnt) {
  const int k = c->map_ngbr_num;
  int rc = knn_async(h, k, (float)c->map_ngbr_max_d);
  if (rc)
    return rc;
  rc = rows_and_wait(h, HtH, Htr, cnt);
  if (h->timing == 1 && rc >= 0) {
    auto el = [&](int a, int b) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]) != cudaSuccess) {
        cudaGetLastError();
        return 0.f;
      }
      return ms;
    };
    const bool ran = h->n_raw > 0;
    h->last_ms[3] = ran ? el(3, 4) : 0.f;
    h->last_ms[4] = ran ? el(4, 5) : 0.f;
    h->last_ms[5] = ran ? el(5, 6) : 0.f;
    h->last_ms[6] = el(6, 8);
  }
  return rc;
}

int ml_linsys(ml_handle *h, const ml_scan_consts *c, const double *P, double *HtH, double *Htr, ml_counters *cnt) {
  if (!h || !P)
    return fail(h, ML_ERR_ARG, "ml_linsys: bad argument");
  RESOLVE(h);
  int rc = check_consts(h, c);
  if (rc)
    return rc;
  if (h->timing)
    cudaEventRecord(h->ev[0], h->stream);
  rc = consts_host(h, c, P);
  if (rc)
    return rc;
  rc = linsys_async_and_wait(h, c, HtH, Htr, cnt);
  if (h->timing && rc >= 0) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev[0], h->ev[8]) == cudaSuccess)
      h->last_ms[7] = ms;
    else
      cudaGetLastError();
    h->last_ms[0] = h->last_ms[1] = h->last_ms[2] = 0.f;
  }
  return rc;
}

} // extern "C"

// The fused front end + kNN of the un-down-sampled one-call path (fused.cuh): one launch.
template <int K, bool VEC4>
static int launch_front_knn(ml_handle *h, const float *raw_src, int stride, int n, float t_rel, float rmin, float rmax, float max_d2,
                            int producers) {
  bool &attr_set = h->fused_attr[K][VEC4 ? 1 : 0];
  if (!attr_set) {
    CK(cudaFuncSetAttribute(k_front_knn<K, VEC4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FK_SMEM));
    attr_set = true;
  }
  const int grid = h->n_sm * 2;
  launch_pdl(h, k_front_knn<K, VEC4>, grid, FK_BLOCK, FK_SMEM, raw_src, stride, n, t_rel, (double)rmin, (double)rmax, h->d_st, h->scan_pts,
             h->scan_orig, h->T_LtoM, h->scan_map, h->lb_state, lb_next_epoch(h), h->d_fctl, std::min(producers, grid),
             (const GridDesc *)h->d_grid, h->q_sorted, h->units, h->units_tag, (const uint32_t *)h->cell_start,
             (const float4 *)h->cell_pts, max_d2, (int32_t *)nullptr, (float *)nullptr, (uint8_t *)nullptr,
             (const float4 *)h->map_pts[h->cur], h->nbr, h->tl_on ? h->tl_knn : (unsigned long long *)nullptr);
  return check_launch(h);
}
template <int K>
static int launch_front_knn_k(ml_handle *h, const float *raw_src, int stride, int n, float t_rel, float rmin, float rmax, float max_d2,
                              int producers) {
  const bool vec4 = stride == 4 && (((uintptr_t)raw_src) & 15u) == 0;
  return vec4 ? launch_front_knn<K, true>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers)
              : launch_front_knn<K, false>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
}
static int scan_front_knn_async(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, const float *T_LtoM,
                                int k, float max_d2) {
  if (n < 0 || (n > 0 && !xyz) || (stride != 3 && stride != 4))
    return fail(h, ML_ERR_ARG, "ml_scan_update: bad argument (stride must be 3 or 4)");
  if (n > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_scan_update: more points than max_scan_points");
  drop_stale_bins(h);
  h->n_raw = n;
  h->rows_valid = false;
  h->vg_ready = false;
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++)
      h->T_LtoM.m[4 * r + c] = T_LtoM[4 * r + c];
  h->have_T = true;
  const float *raw_src = nullptr;
  bool over_pcie = false;
  if (int rc = resolve_input(h, xyz, stride, n, &raw_src, &over_pcie))
    return rc;
  // Producers. From HBM every tile has its own producer and the front end is over in one wave. Over PCIe (the kernel reads
  // pinned host memory itself) a few producers walk the tiles in order, each requesting its next tile while it groups the
  // current one: 32 x 32 KB in flight keep the link full (bandwidth-delay product ~100 KB), the early tiles complete early and
  // the search overlaps the transfer. (A chunked copy-engine transfer with a stream memory operation behind each chunk was
  // measured instead: 2 / 4 / 8 chunks = 136 / 152 / 191 us e2e against 132 us zero-copy - ~5 us per (copy, flag) pair.)
  const int ntiles = cdiv(n, RF_TILE);
  int producers = over_pcie ? 32 : ntiles;
  if (h->fused_producers > 0)
    producers = h->fused_producers;
  producers = std::max(1, std::min(producers, ntiles));
  switch (k) {
  case 3: return launch_front_knn_k<3>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 4: return launch_front_knn_k<4>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 5: return launch_front_knn_k<5>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 6: return launch_front_knn_k<6>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 7: return launch_front_knn_k<7>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 8: return launch_front_knn_k<8>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  case 9: return launch_front_knn_k<9>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  default: return launch_front_knn_k<10>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2, producers);
  }
}

// Arm the parts a k_scan_tile launch will fill (see RR_ARMED_F64 in rows.cuh). rows_host_tail re-arms a part as it consumes it, so in
// steady state this finds everything armed and does nothing.
static void arm_parts(ml_handle *h, int nb, int nred) {
  if (h->parts_armed >= nb && h->parts_armed_nred == nred)
    return;
  for (int b = 0; b < nb; b++) {
    RRPart &p = h->h_parts[b];
    for (int e = 0; e < nred; e++)
      memcpy(&p.v[e], &RR_ARMED_F64, sizeof(double));
    for (int q = 0; q < 5; q++)
      p.c[q] = RR_ARMED_I32;
    p.n_scan = RR_ARMED_I32;
  }
  h->parts_armed = nb;
  h->parts_armed_nred = nred;
}

// ---- the one-kernel path (tile.cuh): a thread owns a ray, a CTA a 448-ray tile down to its share of the reduced system ---------
template <int K, int NB, bool VEC4>
static int launch_scan_tile(ml_handle *h, const float *raw_src, int stride, int n, float t_rel, float rmin, float rmax, float max_d2) {
  using LY = STLayout<NB>;
  int &cap = h->tile_grid[K][NB == 5 ? 1 : 0][VEC4 ? 1 : 0];
  if (!cap) {
    CK(cudaFuncSetAttribute(k_scan_tile<K, NB, VEC4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LY::BYTES));
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_scan_tile<K, NB, VEC4>, ST_BLOCK, LY::BYTES));
    if (per_sm < 1)
      return fail(h, ML_ERR_CUDA, "ml_scan_update: k_scan_tile does not fit an SM");
    // every CTA must be resident: tiles are owned statically and a tile waits for the counts of the tiles before it
    cap = std::min(per_sm * h->n_sm, ML_NPART_MAX);
  }
  const int grid = std::max(1, std::min(cdiv(n, ST_TILE), cap));
  h->parts_n = grid;
  arm_parts(h, grid, LY::NRED);
  launch_pdl(h, k_scan_tile<K, NB, VEC4>, grid, ST_BLOCK, LY::BYTES, h->rowc, raw_src, stride, n, t_rel, (double)rmin, (double)rmax,
             h->d_st, h->scan_pts, h->scan_orig, h->T_LtoM, h->scan_map, h->lb_state, lb_next_epoch(h), (const GridDesc *)h->d_grid,
             (const uint32_t *)h->cell_start, (const float4 *)h->cell_pts, (const float4 *)h->map_pts[h->cur], max_d2,
             h->tile_keep_nbr ? h->nbr : (float4 *)nullptr, h->h_parts,
             h->tl_on ? h->tl_knn : (unsigned long long *)nullptr);
  return check_launch(h);
}
template <int K>
static int launch_scan_tile_k(ml_handle *h, const float *raw_src, int stride, int n, float t_rel, float rmin, float rmax, float max_d2) {
  const bool vec4 = stride == 4 && (((uintptr_t)raw_src) & 15u) == 0;
  if (h->consts.calib_col >= 0)
    return vec4 ? launch_scan_tile<K, 5, true>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2)
                : launch_scan_tile<K, 5, false>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  return vec4 ? launch_scan_tile<K, 3, true>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2)
              : launch_scan_tile<K, 3, false>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
}
static int launch_scan_tile_any(ml_handle *h, int k, const float *raw_src, int stride, int n, float t_rel, float rmin, float rmax, float max_d2) {
  switch (k) {
  case 3: return launch_scan_tile_k<3>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 4: return launch_scan_tile_k<4>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 5: return launch_scan_tile_k<5>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 6: return launch_scan_tile_k<6>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 7: return launch_scan_tile_k<7>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 8: return launch_scan_tile_k<8>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  case 9: return launch_scan_tile_k<9>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  default: return launch_scan_tile_k<10>(h, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
  }
}
// consts_host must have been called (the kernel takes the row constants as a parameter)
static int scan_tile_async(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, const float *T_LtoM,
                           int k, float max_d2) {
  if (n < 0 || (n > 0 && !xyz) || (stride != 3 && stride != 4))
    return fail(h, ML_ERR_ARG, "ml_scan_update: bad argument (stride must be 3 or 4)");
  if (n > h->cap_scan)
    return fail(h, ML_ERR_CAPACITY, "ml_scan_update: more points than max_scan_points");
  drop_stale_bins(h);
  h->n_raw = n;
  h->rows_valid = false;
  h->vg_ready = false;
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 4; c++)
      h->T_LtoM.m[4 * r + c] = T_LtoM[4 * r + c];
  h->have_T = true;
  const float *raw_src = nullptr;
  bool over_pcie = false;
  if (int rc = resolve_input(h, xyz, stride, n, &raw_src, &over_pcie))
    return rc;
  return launch_scan_tile_any(h, k, raw_src, stride, n, t_rel, rmin, rmax, max_d2);
}
// The parts of the launch in flight are consumed by rows_host_tail as their words arrive in pinned memory, so the calling thread
// pays no stream synchronisation (measured: 12 us per call). MINS_TILE_POLL=0 or event timing: synchronise first, then read.
static int scan_tile_finish(ml_handle *h, double *HtH, double *Htr, ml_counters *cnt) {
  if (!h->tile_poll || h->timing) {
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
  }
  return rows_host_tail(h, HtH, Htr, cnt, true);
}

extern "C" {

int ml_scan_update(ml_handle *h, const float *xyz, int stride, int n, float t_rel, float rmin, float rmax, float leaf,
                   const float T_LtoM[16], const ml_scan_consts *c, const double *P, double *HtH, double *Htr, ml_counters *cnt) {
  if (!h || !P || !T_LtoM)
    return fail(h, ML_ERR_ARG, "ml_scan_update: bad argument");
  RESOLVE(h);
  int rc = check_consts(h, c);
  if (rc)
    return rc;
  if (!h->map_init)
    return fail(h, ML_SOFT_NO_MAP, "ml_scan_update: map not initialised");
  if (h->timing)
    cudaEventRecord(h->ev[0], h->stream);
  const bool ds = leaf > 0.f;
  auto el = [&](int a, int b) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev[a], h->ev[b]) != cudaSuccess) {
      cudaGetLastError();
      return 0.f;
    }
    return ms;
  };
  // The one-kernel path pays when a query's walk is at most three levels deep (one thread walks one cube; the level count is that
  // of the kernels: the squared distance is compared with the UNSQUARED option, LidarHelper.cpp:284) and the scan fills the device
  // (a tile per CTA). Measured otherwise: KAIST options (map_ngbr_max_d 10 m = 7 levels on a thinned map) 241 us against
  // 109 us for the grouped walk; a 32 000-ray scan (72 tiles on 148 SMs) 108 us against 93 us. MINS_SCAN_TILE=2 forces it (tests).
  const bool tile_fits =
      h->tile_force || ((int)std::ceil(std::sqrt(c->map_ngbr_max_d) / h->grid_cell * (1.0 + 1e-5)) <= 3 && n >= ST_TILE * h->n_sm);
  if (!ds && h->use_tile && tile_fits && !h->stats_on && h->use_fused_front && n > 0) {
    // ONE launch (tile.cuh). The row constants are a kernel parameter, so the host folds them first (~2 us).
    const bool hp = h->host_prof;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto us = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::micro>(b - a).count();
    };
    const auto t0 = hp ? now() : std::chrono::steady_clock::time_point();
    rc = consts_host(h, c, P);
    if (rc)
      return rc;
    const auto t1 = hp ? now() : t0;
    rc = scan_tile_async(h, xyz, stride, n, t_rel, rmin, rmax, T_LtoM, c->map_ngbr_num, (float)c->map_ngbr_max_d);
    if (rc)
      return rc;
    if (h->timing)
      cudaEventRecord(h->ev[8], h->stream);
    const auto t2 = hp ? now() : t0;
    h->rows_valid = true; // ml_point_rows re-runs the export instantiation of the rows kernel on `nbr` ...
    h->rows_need_knn = !h->tile_keep_nbr; // ... after the multi-call search has produced it, if this kernel did not
    h->rows_exported = false;
    h->rows_k = c->map_ngbr_num;
    rc = scan_tile_finish(h, HtH, Htr, cnt);
    if (hp) { // MINS_HOST_PROF=1: where the calling thread spends the call (printed by ml_destroy)
      h->host_prof_us[0] += us(t0, t1), h->host_prof_us[1] += us(t1, t2), h->host_prof_us[2] += us(t2, now());
      h->host_prof_n++;
    }
    if (h->timing && rc >= 0) {
      for (int q = 0; q < 7; q++)
        h->last_ms[q] = 0.f;
      h->last_ms[7] = el(0, 8);
      h->last_ms[4] = h->last_ms[7]; // reported under the search stage: one kernel
    }
    return rc;
  }
  if (!ds && h->use_fused_knn && h->use_fused_front && n > 0) {
    // two launches: front end + kNN (fused.cuh), rows + reduction (rows.cuh). The host folds the scan constants while the first
    // one runs; nothing sits between the two kernels in the stream (programmatic dependent launch).
    rc = scan_front_knn_async(h, xyz, stride, n, t_rel, rmin, rmax, T_LtoM, c->map_ngbr_num, (float)c->map_ngbr_max_d);
    if (rc)
      return rc;
    rc = consts_host(h, c, P);
    if (rc)
      return rc;
    if (h->timing == 1)
      cudaEventRecord(h->ev[5], h->stream);
    rc = rows_and_wait(h, HtH, Htr, cnt);
    if (h->timing && rc >= 0) {
      for (int q = 0; q < 7; q++)
        h->last_ms[q] = 0.f;
      if (h->timing == 1) {
        h->last_ms[4] = el(0, 5); // front end + cell grouping + kNN: one kernel
        h->last_ms[5] = el(5, 6);
      }
      h->last_ms[7] = el(0, 8);
    }
    return rc;
  }
  // without down-sampling the compaction pass also writes pointcloud_in_map
  rc = scan_upload_async(h, xyz, stride, n, t_rel, rmin, rmax, ds ? nullptr : T_LtoM, ds ? leaf : 0.f);
  if (rc)
    return rc;
  // the constants are folded after the scan is enqueued, so that the GPU already works meanwhile
  rc = consts_host(h, c, P);
  if (rc)
    return rc;
  if (h->timing == 1)
    cudaEventRecord(h->ev[1], h->stream);
  if (ds) {
    rc = scan_downsample_async(h, leaf, T_LtoM);
    if (rc)
      return rc;
  }
  if (h->timing == 1)
    cudaEventRecord(h->ev[2], h->stream);
  if (ds && !h->bin_done) { // not fused (MINS_NO_FUSED_FRONT=1): separate transform, binning inside the kNN stage
    rc = scan_transform_async(h, T_LtoM);
    if (rc)
      return rc;
  }
  if (h->timing == 1)
    cudaEventRecord(h->ev[9], h->stream);
  rc = linsys_async_and_wait(h, c, HtH, Htr, cnt);
  if (h->timing && rc >= 0) {
    if (h->timing == 1) {
      h->last_ms[0] = el(0, 1);
      h->last_ms[1] = el(1, 2);
      h->last_ms[2] = el(2, 9);
    } else {
      for (int q = 0; q < 7; q++)
        h->last_ms[q] = 0.f;
    }
    h->last_ms[7] = el(0, 8);
  }
  return rc;
}

int ml_point_rows(ml_handle *h, ml_point_row *rows, double *h_rows, int cap, int *n) {
  if (!h)
    return ML_ERR_ARG;
  RESOLVE(h);
  if (!h->rows_valid)
    return fail(h, ML_ERR_ARG, "ml_point_rows: no ml_linsys result for the current scan and map (the export must follow "
                               "ml_linsys / ml_scan_update directly)");
  if (h->rows_need_knn) {
    // the one-kernel path kept no neighbour coordinates: the multi-call search on the resident scan and map writes them (the
    // same neighbours: every walk has one arithmetic and one order), then the export runs as after ml_linsys
    if (int rc = knn_async(h, h->rows_k, (float)h->consts.map_ngbr_max_d))
      return rc;
    h->rows_need_knn = false;
  }
  const int cnt = h->last_n_scan;
  const int hx = h->consts.total_hx;
  const int m = std::min(cap, cnt);
  if (n)
    *n = cnt;
  if (m <= 0)
    return ML_OK;
  if (!h->rows_exported && (rows || h_rows)) {
    // the timed path keeps the per-point rows in shared memory only; re-run the same kernel with the export switched on
    // (identical inputs are still resident, the result block is reproduced bit for bit)
    int rc = launch_rows_k(h, h->rows_k, h->rows);
    if (rc)
      return rc;
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    h->rows_exported = true;
  }
  if (rows) {
    std::vector<PointRow> tmp((size_t)m);
    CK(cudaMemcpyAsync(tmp.data(), h->rows, sizeof(PointRow) * (size_t)m, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    for (int i = 0; i < m; i++) {
      ml_point_row &o = rows[i];
      const PointRow &r = tmp[i];
      o.status = r.status;
      o.pad = 0;
      memcpy(o.plane, r.plane, sizeof(o.plane));
      o.res = r.res, o.chi2 = r.chi2, o.w = r.w, o.vr = r.vr, o.raw_n = r.raw_n;
    }
  }
  if (h_rows) {
    const size_t need = (size_t)cnt * hx;
    if (need > h->hrows_cap) { // debug export only: grown on demand and kept
      if (h->d_hrows)
        cudaFree(h->d_hrows);
      h->d_hrows = nullptr, h->hrows_cap = 0;
      CK(cudaMalloc(&h->d_hrows, sizeof(double) * need));
      h->hrows_cap = need;
    }
    CK(cudaMemcpyAsync(h->d_T, h->hT, sizeof(double) * ML_G * ML_MAX_HX, cudaMemcpyHostToDevice, h->stream));
    LAUNCH(k_expand_h, std::min(148 * 8, cdiv((long long)cnt * hx, 256)), 256, h->d_st, h->rows, (const double *)h->d_T, hx, h->d_hrows);
    CK(cudaMemcpyAsync(h_rows, h->d_hrows, sizeof(double) * (size_t)m * hx, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
  }
  return ML_OK;
}

// Debug/profiling aid (not part of the reference's surface): arm (on = 1) or disarm per-CTA %globaltimer stamps in the kNN and
// rows kernels; with out != NULL copy the stamps of the last scan: out[0 .. cap) receives 2 * ML_TL_MAX_CTAS * 8 values
// ([kernel: 0 kNN, 1 rows][cta][slot], ns, 0 = not reached). Returns the number of values written or < 0.
int ml_debug_timeline(ml_handle *h, int on, unsigned long long *out, int cap) {
  if (!h)
    return ML_ERR_ARG;
  use_device(h);
  const size_t per = (size_t)ML_TL_MAX_CTAS * ML_TL_SLOTS;
  if (!h->tl_rows) {
    CK(cudaMalloc(&h->tl_rows, sizeof(unsigned long long) * per * 2));
    h->tl_knn = h->tl_rows + per;
    CK(cudaMemset(h->tl_rows, 0, sizeof(unsigned long long) * per * 2));
  }
  int written = 0;
  if (out && cap > 0) {
    CK(cudaStreamSynchronize(h->stream));
    const size_t m = std::min((size_t)cap, per * 2);
    std::vector<unsigned long long> tmp(per * 2);
    CK(cudaMemcpy(tmp.data(), h->tl_rows, sizeof(unsigned long long) * per * 2, cudaMemcpyDeviceToHost));
    // device order is [rows | knn]; hand out [knn | rows]
    for (size_t i = 0; i < m; i++)
      out[i] = i < per ? tmp[per + i] : tmp[i - per];
    written = (int)m;
    CK(cudaMemset(h->tl_rows, 0, sizeof(unsigned long long) * per * 2));
  }
  h->tl_on = on != 0;
  return written;
}

void ml_abi_sizes(int out[4]) {
  out[0] = (int)sizeof(ml_scan_consts);
  out[1] = (int)sizeof(ml_point_row);
  out[2] = (int)sizeof(ml_counters);
  out[3] = (int)sizeof(ml_config);
}

int ml_set_stats(ml_handle *h, int on) {
  if (!h)
    return ML_ERR_ARG;
  h->stats_on = on != 0;
  return ML_OK;
}

int ml_set_timing(ml_handle *h, int on) {
  if (!h)
    return ML_ERR_ARG;
  h->timing = on < 0 ? 0 : (on > 2 ? 2 : on);
  return ML_OK;
}

int ml_last_timing(ml_handle *h, float ms[8]) {
  if (!h || !ms)
    return ML_ERR_ARG;
  memcpy(ms, h->last_ms, sizeof(h->last_ms));
  return ML_OK;
}

// ---------------------------------------------------------------------------------------------------
// Host tail: information form -> square-root form by Cholesky with diagonal pivoting (rank revealing), the
// stand-in for StateHelper::measurement_compress_inplace (StateHelper.cpp:631-643): Rc^T Rc = HtH, Rc^T rc = Htr.
// ---------------------------------------------------------------------------------------------------
int ml_info_to_sqrt(const double *HtH, const double *Htr, int n, double *Rc, double *rc) {
  if (!HtH || !Htr || !Rc || !rc || n < 1 || n > ML_MAX_HX)
    return ML_ERR_ARG;
  std::vector<double> A(HtH, HtH + (size_t)n * n), L((size_t)n * n, 0.0);
  std::vector<int> perm(n);
  for (int i = 0; i < n; i++)
    perm[i] = i;
  double maxd = 0;
  for (int i = 0; i < n; i++)
    maxd = std::max(maxd, A[(size_t)i * n + i]);
  const double tol = maxd * (double)n * 2.220446049250313e-16 * 16.0;
  int rank = 0;
  for (int k = 0; k < n; k++) {
    int p = k;
    double best = -1;
    for (int i = k; i < n; i++) {
      double d = A[(size_t)perm[i] * n + perm[i]];
      for (int j = 0; j < k; j++)
        d -= L[(size_t)perm[i] * n + j] * L[(size_t)perm[i] * n + j];
      if (d > best)
        best = d, p = i;
    }
    if (!(best > tol))
      break;
    std::swap(perm[k], perm[p]);
    const int pk = perm[k];
    const double lkk = std::sqrt(best);
    L[(size_t)pk * n + k] = lkk;
    for (int i = k + 1; i < n; i++) {
      const int pi = perm[i];
      double s = A[(size_t)pi * n + pk];
      for (int j = 0; j < k; j++)
        s -= L[(size_t)pi * n + j] * L[(size_t)pk * n + j];
      L[(size_t)pi * n + k] = s / lkk;
    }
    rank++;
  }
  // Rc (rank x n): Rc[k][col] = L[col][k];  rc: forward solve L[perm[0..rank)][0..rank) rc = Htr[perm]
  for (int k = 0; k < n; k++)
    for (int c = 0; c < n; c++)
      Rc[(size_t)k * n + c] = k < rank ? L[(size_t)c * n + k] : 0.0;
  for (int k = 0; k < n; k++)
    rc[k] = 0.0;
  for (int k = 0; k < rank; k++) {
    double s = Htr[perm[k]];
    for (int j = 0; j < k; j++)
      s -= L[(size_t)perm[k] * n + j] * rc[j];
    rc[k] = s / L[(size_t)perm[k] * n + k];
  }
  return rank;
}

} // extern "C"

// ------------------------------------------------------------------------------------------------
// ml_range_filter_static: the range filter of ml_scan_upload (LiDARData's constructor as a stable two-launch compaction,
// k_range_count + k_range_compact in scan.cuh) for callers that cannot allocate - CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc. No handle: the per-tile counts that the handle path keeps in its `tile_sums` buffer live in
// a file-scope device array of fixed size; inputs and outputs are the caller's. No counterpart in the
// reference.
//   * the array holds one count per 2048-point tile: ML_STATIC_MAX_POINTS / 2048 words = 8 KB on every GPU that loads the
//     library, until the process exits;
//   * in HOST code `g_static_tile_sums` is only a stand-in for the symbol: its device address comes from
//     cudaGetSymbolAddress, once per device with that device current (static_scratch below), and is what the kernels get;
//   * every call on a device shares the array and finds the previous call's counts in it. Nothing has to start at zero: the
//     count kernel overwrites exactly the words [0, ceil(n / 2048)) that the compaction kernel of the same call reads. Two calls
//     that overlap on the GPU (different streams, or concurrent graph launches) race on it.
// ------------------------------------------------------------------------------------------------
namespace ml {
constexpr int STATIC_MAX_TILES = ML_STATIC_MAX_POINTS / RF_TILE;
static __device__ __align__(256) unsigned int g_static_tile_sums[STATIC_MAX_TILES];

template <bool VEC4>
__global__ void __launch_bounds__(RF_BLOCK) k_range_count_static(const float *__restrict__ raw, int stride, int n, double rmin, double rmax,
                                                                 unsigned int *__restrict__ tile_sums) {
  __shared__ uint32_t wsum[RF_BLOCK / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int wbase = blockIdx.x * RF_TILE + wid * (RF_TILE / (RF_BLOCK / 32));
  uint32_t cnt = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int i = wbase + r * 32 + lane;
    const bool k = i < n && range_keep(load_raw<VEC4>(raw, stride, i), rmin, rmax);
    cnt += __popc(__ballot_sync(0xffffffffu, k));
  }
  if (lane == 0)
    wsum[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
#pragma unroll
    for (int w = 0; w < RF_BLOCK / 32; w++)
      s += wsum[w];
    tile_sums[blockIdx.x] = s;
  }
}

// the body of k_range_compact with one output cloud and the count delivered to the caller's device word
template <bool VEC4>
__global__ void __launch_bounds__(RF_BLOCK)
    k_range_compact_static(const float *__restrict__ raw, int stride, int n, float t_rel, double rmin, double rmax,
                           const unsigned int *__restrict__ tile_sums, float4 *__restrict__ out, int *__restrict__ n_kept) {
  __shared__ uint32_t wsum[RF_BLOCK / 32];
  __shared__ uint32_t tile_prefix;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (wid == 0) { // counts of the tiles before this one, fixed order
    uint32_t s = 0;
    for (int b = lane; b < (int)blockIdx.x; b += 32)
      s += tile_sums[b];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0)
      tile_prefix = s;
  }
  const int wbase = blockIdx.x * RF_TILE + wid * (RF_TILE / (RF_BLOCK / 32));
  float4 pt[8];
  uint32_t bal[8];
  uint32_t cnt = 0;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int i = wbase + r * 32 + lane;
    pt[r] = i < n ? load_raw<VEC4>(raw, stride, i) : make_float4(0.f, 0.f, 0.f, 0.f);
    const bool k = i < n && range_keep(pt[r], rmin, rmax);
    bal[r] = __ballot_sync(0xffffffffu, k);
    cnt += __popc(bal[r]);
  }
  if (lane == 0)
    wsum[wid] = cnt;
  __syncthreads();
  uint32_t pos = tile_prefix;
  for (int w = 0; w < wid; w++)
    pos += wsum[w];
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
    uint32_t tot = tile_prefix;
#pragma unroll
    for (int w = 0; w < RF_BLOCK / 32; w++)
      tot += wsum[w];
    *n_kept = (int)tot;
  }
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < 8; r++) {
    if ((bal[r] >> lane) & 1u)
      out[pos + __popc(bal[r] & lt)] = make_float4(pt[r].x, pt[r].y, pt[r].z, t_rel);
    pos += __popc(bal[r]);
  }
}
} // namespace ml

// device address of the scratch array on the CURRENT device; looked up once per device
static int static_scratch(unsigned int **out) {
  ml_handle *h = nullptr; // CK reports through ml_global_error
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static unsigned int *addr[MAX_DEV] = {};
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEV)
    return fail(nullptr, ML_ERR_ARG, "ml_range_filter_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    unsigned int *p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, ml::g_static_tile_sums));
    if (!p)
      return fail(nullptr, ML_ERR_CUDA, "ml_range_filter_static: cudaGetSymbolAddress returned no address");
    addr[dev] = p;
  }
  *out = addr[dev];
  return ML_OK;
}

extern "C" int ml_range_filter_static(const float *xyz_dev, int stride, int n, float t_rel, float rmin, float rmax, float *xyzi_out_dev,
                                      int *n_kept_dev, void *stream) {
  ml_handle *h = nullptr;
  if (n < 0 || (stride != 3 && stride != 4) || !n_kept_dev || (n > 0 && (!xyz_dev || !xyzi_out_dev)))
    return fail(nullptr, ML_ERR_ARG, "ml_range_filter_static: bad argument (stride must be 3 or 4, pointers must not be NULL)");
  if (n > ML_STATIC_MAX_POINTS)
    return fail(nullptr, ML_ERR_CAPACITY, "ml_range_filter_static: n exceeds ML_STATIC_MAX_POINTS (4194304 points): the scratch array "
                                          "is static; use ml_scan_upload for larger scans");
  if ((((uintptr_t)xyzi_out_dev) & 15u) != 0)
    return fail(nullptr, ML_ERR_ARG, "ml_range_filter_static: xyzi_out_dev must be 16-byte aligned");
  unsigned int *sums = nullptr;
  if (int rc = static_scratch(&sums))
    return rc;
  cudaStream_t s = (cudaStream_t)stream;
  const int tiles = std::max(1, cdiv(n, RF_TILE)); // n == 0: one CTA that keeps nothing and writes n_kept = 0
  const bool vec4 = stride == 4 && (((uintptr_t)xyz_dev) & 15u) == 0;
  if (vec4) {
    k_range_count_static<true><<<tiles, RF_BLOCK, 0, s>>>(xyz_dev, stride, n, (double)rmin, (double)rmax, sums);
    k_range_compact_static<true><<<tiles, RF_BLOCK, 0, s>>>(xyz_dev, stride, n, t_rel, (double)rmin, (double)rmax, sums,
                                                            (float4 *)xyzi_out_dev, n_kept_dev);
  } else {
    k_range_count_static<false><<<tiles, RF_BLOCK, 0, s>>>(xyz_dev, stride, n, (double)rmin, (double)rmax, sums);
    k_range_compact_static<false><<<tiles, RF_BLOCK, 0, s>>>(xyz_dev, stride, n, t_rel, (double)rmin, (double)rmax, sums,
                                                             (float4 *)xyzi_out_dev, n_kept_dev);
  }
  CK(cudaGetLastError());
  return ML_OK;
}

#include "icp_host.inl"
