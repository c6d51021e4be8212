#pragma once
#include <string>
#include <ros/time.h>
namespace std_msgs { struct Header { unsigned seq; ros::Time stamp; std::string frame_id; Header() : seq(0) {} }; }
