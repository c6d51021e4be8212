#pragma once
namespace geometry_msgs { struct Vector3 { double x, y, z; Vector3() : x(0), y(0), z(0) {} }; }
