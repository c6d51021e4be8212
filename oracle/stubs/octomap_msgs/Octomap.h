#pragma once
#include <std_msgs/Header.h>
namespace octomap_msgs { struct Octomap { std_msgs::Header header; int binary; int id; double resolution; }; }
