// TEST INFRASTRUCTURE ONLY. Minimal stand-in for <pcl/point_types.h> so that the reference's own
// mins/src/update/lidar/ikd_Tree.cpp compiles UNMODIFIED where it lies under /root/reference
// (PCL itself is not installed in this image). Only the three point structs ikd_Tree.cpp instantiates
// are declared; layout follows PCL's 16-byte aligned float structs (xyz+pad, intensity in the 2nd quad).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
namespace Eigen {
template <class T> using aligned_allocator = std::allocator<T>;
}
namespace pcl {
struct alignas(16) PointXYZ {
  float x = 0, y = 0, z = 0, pad = 1;
};
struct alignas(16) PointXYZI {
  float x = 0, y = 0, z = 0, pad = 1;
  float intensity = 0;
  float p2[3] = {0, 0, 0};
};
struct alignas(16) PointXYZINormal {
  float x = 0, y = 0, z = 0, pad = 1;
  float nx = 0, ny = 0, nz = 0, p1 = 0;
  float intensity = 0, curvature = 0;
  float p2[2] = {0, 0};
};
} // namespace pcl
