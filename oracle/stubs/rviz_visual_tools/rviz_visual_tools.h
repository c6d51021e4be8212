#pragma once
// headless stand-in: every drawing call is a no-op (no graphics stack on this box)
#include <memory>
#include <string>
namespace rviz_visual_tools {
enum colors { RED, GREEN, BLUE, PURPLE, YELLOW, ORANGE, CYAN, DARK_GREY };
enum scales { LARGE };
struct RvizVisualTools {
  RvizVisualTools(const std::string&, const std::string&) {}
  template <class A, class B> void publishLine(const A&, const B&, colors, scales = LARGE) {}
  void trigger() {}
};
typedef std::shared_ptr<RvizVisualTools> RvizVisualToolsPtr;
}
