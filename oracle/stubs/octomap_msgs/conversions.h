#pragma once
#include <octomap_msgs/Octomap.h>
namespace octomap_msgs { template <class T> inline bool fullMapToMsg(const T&, Octomap&) { return false; } }
