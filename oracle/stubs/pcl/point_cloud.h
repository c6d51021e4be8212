#pragma once
#include <memory>
#include <vector>
#include <pcl/point_types.h>
namespace pcl {
template <class P> struct PointCloud {
  typedef std::shared_ptr<PointCloud<P> > Ptr;
  std::vector<P> points;
  void push_back(const P& p) { points.push_back(p); }
  size_t size() const { return points.size(); }
  PointCloud& operator+=(const PointCloud& o) { points.insert(points.end(), o.points.begin(), o.points.end()); return *this; }
};
}
