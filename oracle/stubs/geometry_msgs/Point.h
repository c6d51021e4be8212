#pragma once
// stand-in for the ROS message type (plain doubles, zero-initialised like genmsg output)
namespace geometry_msgs { struct Point { double x, y, z; Point() : x(0), y(0), z(0) {} }; }
