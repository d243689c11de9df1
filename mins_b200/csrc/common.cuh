// Shared device/host declarations of the B200 LiDAR-update library (sm_100a only).
// Reference citations are relative to /root/reference.
#pragma once
#include "../../include/mins_lidar.h"
#include <cuda_runtime.h>
#include <stdint.h>

namespace ml {

// ------------------------------------------------------------------------------------------------
// fp32 arithmetic that must round like the reference's baseline x86-64 build (no FMA contraction,
// mins/CMakeLists.txt:15): explicit _rn intrinsics are never fused by nvcc.
// ------------------------------------------------------------------------------------------------
// KD_TREE::calc_dist, ikd_Tree.cpp:1406-1410: (dx*dx + dy*dy) + dz*dz
__device__ __forceinline__ float dist2_ref(float ax, float ay, float az, float bx, float by, float bz) {
  float dx = __fsub_rn(ax, bx), dy = __fsub_rn(ay, by), dz = __fsub_rn(az, bz);
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// Eigen Vector3f::norm() as evaluated for a fixed-size 3-vector: sqrt(x*x + (y*y + z*z))
__device__ __forceinline__ float norm3_ref(float x, float y, float z) {
  return __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z))));
}

// order-preserving float <-> uint mapping for atomicMin/Max on floats
__device__ __forceinline__ uint32_t f2ord(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// ---- pcl::transformPointCloud(cloud, cloud_in_map, Matrix4f), LidarHelper.cpp:192,203 -----------------
// row-major 3x4 fp32 rigid transform. PCL 1.10 SSE path: m0*x + (m1*y + (m2*z + m3)) per row, fp32, no FMA.
struct Tf32 {
  float m[12];
};
__device__ __forceinline__ float4 tf_apply(const Tf32 &T, float4 p) {
  float4 q;
  q.x = __fadd_rn(__fmul_rn(T.m[0], p.x), __fadd_rn(__fmul_rn(T.m[1], p.y), __fadd_rn(__fmul_rn(T.m[2], p.z), T.m[3])));
  q.y = __fadd_rn(__fmul_rn(T.m[4], p.x), __fadd_rn(__fmul_rn(T.m[5], p.y), __fadd_rn(__fmul_rn(T.m[6], p.z), T.m[7])));
  q.z = __fadd_rn(__fmul_rn(T.m[8], p.x), __fadd_rn(__fmul_rn(T.m[9], p.y), __fadd_rn(__fmul_rn(T.m[10], p.z), T.m[11])));
  q.w = p.w;
  return q;
}

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch (sm_90+): the kernels of the per-scan pipeline are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, call pdl_launch() first thing (the next kernel's CTAs may become
// resident as soon as every CTA of this one has started) and pdl_wait() before touching global memory (returns when the
// preceding kernel has completed and flushed). Only launch latency and CTA scheduling overlap; the data dependencies are
// those of plain stream order. Without the launch attribute both instructions are no-ops.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_enter() {
  pdl_launch();
  pdl_wait();
}

// Debug timeline (ml_debug_timeline): thread 0 of every CTA stamps %globaltimer [ns] at up to ML_TL_SLOTS points of a kernel.
// tl == NULL (always, unless the debug entry point armed it): one predicated-off store per mark.
#define ML_TL_SLOTS 8
__device__ __forceinline__ void tl_mark(unsigned long long *tl, int slot) {
  if (tl && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tl[(size_t)blockIdx.x * ML_TL_SLOTS + slot] = t;
  }
}

// ------------------------------------------------------------------------------------------------
// Dense kNN grid over the map's bounding box. Cell id = (z*dimy + y)*dimx + x (x fastest, so the cells
// of one x-run are one contiguous range of cell_pts).
// ------------------------------------------------------------------------------------------------
struct GridDesc {
  double ox, oy, oz; // origin (lower corner of cell 0,0,0)
  double cell, inv_cell;
  int dimx, dimy, dimz;
  int ncells;
  int n_pts; // map points binned
  int valid;
};

__device__ __forceinline__ int cell_coord(double x, double o, double inv) { return (int)floor((x - o) * inv); }

// Device-resident state of the current scan (counts produced on the device, read by later kernels).
struct ScanState {
  int n_raw;  // points uploaded
  int n_orig; // after the range filter  (pointcloud_original)
  int n_scan; // after down-sampling      (pointcloud / pointcloud_in_map)
  int vox_n;  // occupied voxels of the last voxel-grid filter
  int vox_passthrough;
  int pad;
  // kNN work-list counters, double-buffered by the parity of the kNN call: call c fills [c & 1] and clears [(c + 1) & 1]
  int n_units[2];   // warp work units
  int n_qsorted[2]; // queries placed in cell order
  int knn_next[2];  // kNN work ticket: units beyond the first one of every warp are handed out dynamically
};

// One warp work unit of the kNN kernel: up to 32 queries that fall into the same grid cell.
struct KnnUnit {
  int qbase, cnt;
  int cx, cy, cz;
};

// Per-point Jacobian direction g (UpdaterLidar.cpp:320-329), folded: the reference's 18 numbers
// [n dpfinM_dI (6) | n dpfinM_dA (6) | n dpfinM_dcalib (6)] contain n^T RGtoM twice with opposite signs (position block of I,
// position block of A), so 15 independent numbers remain:
//   [ I rotation (3) | n^T RGtoM (3; multiplies dTdxI position rows MINUS dTdxA position rows) | A rotation (3) | calib (6) ]
// and only the first 9 are non-zero without extrinsic calibration. The per-scan constants derived from ml_scan_consts + P
// (T: h = g^T T; Q = T P T^T) are folded on the host: RowConsts (rows.cuh), consts_host (api.cu).
#define ML_G 15

// Per-point result of point_row (k_rows_reduce), consumed by the fixed-order reduction.
struct PointRow {
  double g[ML_G]; // whitened: g / raw_noise
  double w;       // ||v||^2
  double vr;      // v^T r'
  double chi2;
  double res;     // whitened scan-point residual
  double raw_n;
  double rss;     // ||r'||^2
  double plane[4];
  int status;
  int pad;
};

#define ML_NPART_MAX 512 // >= CTAs of the kernels that return one reduced system each (k_rows_reduce: one per SM; k_scan_tile: two)
#define ML_GSYM (ML_G * (ML_G + 1) / 2) // 120

} // namespace ml
