#pragma once
// volumetric_mapping::OctomapManager as the planner sees it (src/distance_heuristic.cpp:108-111):
// a thin forwarder to the restated OctomapWorld in oracle/octomap_restated.hpp.
#include <Eigen/Dense>
#include "octomap_restated.hpp"
namespace volumetric_mapping {
class OctomapManager {
 public:
  enum CellStatus { kFree = 0, kOccupied = 1, kUnknown = 2 };
  explicit OctomapManager(orc::OctomapWorld* w) : world(w), calls(0) {}
  CellStatus getLineStatusBoundingBox(const Eigen::Vector3d& start, const Eigen::Vector3d& end,
                                      const Eigen::Vector3d& box) const {
    ++calls;
    return (CellStatus)world->getLineStatusBoundingBox(start.v, end.v, box.v);
  }
  orc::OctomapWorld* world;
  mutable long calls;
};
}
