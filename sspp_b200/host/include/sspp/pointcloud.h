// Minimal stand-ins for the two PCL types that appear in the Heuristic interface
// (pcl::PointXYZ, pcl::PointCloud<pcl::PointXYZ>), so the interface keeps its shape without PCL.
// Inside a PCL build define SSPP_WITH_PCL and the real headers are used.
#pragma once
#ifdef SSPP_WITH_PCL
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#else
#include <memory>
#include <vector>
namespace pcl {
struct PointXYZ { float x = 0, y = 0, z = 0; PointXYZ() {} PointXYZ(float a, float b, float c) : x(a), y(b), z(c) {} };
template <class P> struct PointCloud {
  typedef std::shared_ptr<PointCloud<P> > Ptr;
  std::vector<P> points;
  void push_back(const P& p) { points.push_back(p); }
  size_t size() const { return points.size(); }
  PointCloud& operator+=(const PointCloud& o) { points.insert(points.end(), o.points.begin(), o.points.end()); return *this; }
};
}  // namespace pcl
#endif
