#pragma once
// headless stand-in for roscpp: publishers swallow messages, parameters keep their defaults
// real roscpp headers drag these in; the planner sources rely on that
#include <unistd.h>
#include <cerrno>
#include <cstring>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <algorithm>
#include <map>
#include <vector>
#include <string>
#include <ros/time.h>
namespace ros {
struct Publisher { template <class M> void publish(const M&) const {} };
struct NodeHandle {
  template <class M> Publisher advertise(const std::string&, int) { return Publisher(); }
  template <class T> bool param(const std::string&, T& v, const T& d) const { v = d; return false; }
  template <class T> bool getParam(const std::string&, T&) const { return false; }
};
inline bool ok() { return false; }
inline void spinOnce() {}
}
#define ROS_INFO(...) do {} while (0)
#define ROS_WARN(...) do {} while (0)
#define ROS_ERROR(...) do {} while (0)
#define ROS_INFO_THROTTLE(...) do {} while (0)
#define ROS_INFO_STREAM(x) do {} while (0)
