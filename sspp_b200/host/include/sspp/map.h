// Map — 2-D boolean grid (reference include/sspp/map.h).  Named by the planner API but read by no
// code path of the reference; kept as a plain value type.
#pragma once
#include <vector>
class Map {
 public:
  int width = 0, height = 0;
  float mapRes = 0;
  std::vector<std::vector<bool> > grid;
  Map() {}
  Map(int w, int h, float res) : width(w), height(h), mapRes(res), grid(w, std::vector<bool>(h, false)) {}
};
