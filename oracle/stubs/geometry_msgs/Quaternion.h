#pragma once
namespace geometry_msgs { struct Quaternion { double x, y, z, w; Quaternion() : x(0), y(0), z(0), w(0) {} }; }
