#pragma once
// placeholder types: SSPP::Node carries these as unused payload members (include/sspp/node.h:51-56)
namespace octomap {
struct OcTree { explicit OcTree(double) {} };
struct Pointcloud {};
}
