#pragma once
#include <chrono>
#include <thread>
namespace ros {
struct Time {
  double t; Time() : t(0) {}
  static Time now() { Time r; r.t = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); return r; }
  double toSec() const { return t; }
};
struct Duration { double d; Duration(double s = 0) : d(s) {} void sleep() const {} double toSec() const { return d; } };
struct Rate { Rate(double) {} void sleep() {} };
}
