#pragma once
// stand-in for pcl::KdTreeFLANN: exact brute-force 1-NN (only PathPlanner::connectToNN uses it)
#include <pcl/point_cloud.h>
namespace pcl {
template <class P> struct KdTreeFLANN {
  typename PointCloud<P>::Ptr cloud;
  void setInputCloud(const typename PointCloud<P>::Ptr& c) { cloud = c; }
  int nearestKSearch(const P& q, int k, std::vector<int>& idx, std::vector<float>& d2) const {
    if (!cloud || cloud->points.empty() || k < 1) return 0;
    int best = 0; float bd = -1;
    for (size_t i = 0; i < cloud->points.size(); ++i) {
      const P& p = cloud->points[i];
      float d = (p.x - q.x) * (p.x - q.x) + (p.y - q.y) * (p.y - q.y) + (p.z - q.z) * (p.z - q.z);
      if (bd < 0 || d < bd) { bd = d; best = (int)i; }
    }
    idx.assign(1, best); d2.assign(1, bd);
    return 1;
  }
};
}
