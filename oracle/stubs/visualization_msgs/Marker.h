#pragma once
#include <vector>
#include <std_msgs/Header.h>
#include <geometry_msgs/Pose.h>
#include <geometry_msgs/Vector3.h>
namespace visualization_msgs { struct Marker { std_msgs::Header header; }; }
