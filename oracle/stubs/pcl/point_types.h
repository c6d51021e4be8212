#pragma once
namespace pcl { struct PointXYZ { float x, y, z; PointXYZ() : x(0), y(0), z(0) {} PointXYZ(float a, float b, float c) : x(a), y(b), z(c) {} }; }
